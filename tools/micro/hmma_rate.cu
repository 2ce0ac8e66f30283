// Issue rate of the legacy warp-level MMA path on sm_100a (B200): mma.sync m16n8k8 TF32 and m16n8k16 BF16, with
// 1..4 resident warps per SM sub-partition and 4 independent accumulator chains per warp, timed with clock64.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/hmma_rate tools/micro/hmma_rate.cu && ./hmma_rate
#include <cstdio>
#include <cuda_runtime.h>

template <bool BF16>
__global__ void rate_kernel(int iters, long long* cycles, float* sink) {
  float c[4][4] = {};
  unsigned a[4] = {0x3f800000u + threadIdx.x, 0x3f810000u, 0x3f820000u, 0x3f830000u};
  unsigned b[2] = {0x3f840000u, 0x3f850000u + threadIdx.x};
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (BF16)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[q][0]), "+f"(c[q][1]), "+f"(c[q][2]), "+f"(c[q][3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
      else
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[q][0]), "+f"(c[q][1]), "+f"(c[q][2]), "+f"(c[q][3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = c[0][0] + c[1][1] + c[2][2] + c[3][3];
}

int main() {
  long long* cyc;
  float* sink;
  cudaMalloc(&cyc, 1024 * sizeof(long long));
  cudaMalloc(&sink, 1024 * 1024 * sizeof(float));
  const int iters = 4096;
  for (int bf16 = 0; bf16 < 2; ++bf16)
    for (int warps = 4; warps <= 16; warps *= 2) {
      for (int rep = 0; rep < 2; ++rep) {
        if (bf16) rate_kernel<true><<<148, warps * 32>>>(iters, cyc, sink);
        else rate_kernel<false><<<148, warps * 32>>>(iters, cyc, sink);
      }
      cudaDeviceSynchronize();
      long long h[148];
      cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
      double mean = 0;
      for (int i = 0; i < 148; ++i) mean += (double)h[i] / 148;
      const double mma_per_smsp = (double)iters * 4 * warps / 4;
      const double fma = bf16 ? 16.0 * 8 * 16 : 16.0 * 8 * 8;
      printf("%s  %2d warps/SM (%d per sub-partition): %.2f cycles per MMA per sub-partition, %.0f FMA/clk/SM  [%s]\n",
             bf16 ? "m16n8k16 bf16" : "m16n8k8  tf32", warps, warps / 4, mean / mma_per_smsp, fma * 4 / (mean / mma_per_smsp),
             cudaGetErrorString(cudaGetLastError()));
    }
  return 0;
}
