// Microbenchmark: legacy warp-level mma.sync / ldmatrix throughput on sm_100a (per SM sub-partition).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/hmma_bench tools/hmma_bench.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
__global__ void k_hmma(float* out, int iters, int chains) {
  float d[4][4] = {};
  uint32_t a = 0x3F803F80u, b0 = 0x3F803F80u + threadIdx.x, b1 = 0x3F003F00u;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if (c < chains)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%4,%4,%4}, {%5,%6}, {%0,%1,%2,%3};"
                     : "+f"(d[c][0]), "+f"(d[c][1]), "+f"(d[c][2]), "+f"(d[c][3]) : "r"(a), "r"(b0), "r"(b1));
    }
  }
  long long t1 = clock64();
  float s = 0; for (int c = 0; c < 4; ++c) s += d[c][0] + d[c][1] + d[c][2] + d[c][3];
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (float)(t1 - t0);
  if (s == 12345.f) out[1] = s;
}
__global__ void k_ldsm(float* out, int iters) {
  __shared__ __align__(16) uint8_t sm[8 * 2048];
  for (int i = threadIdx.x; i < 8 * 2048 / 4; i += blockDim.x) ((uint32_t*)sm)[i] = i;
  __syncthreads();
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  uint32_t base = (uint32_t)__cvta_generic_to_shared(sm + w * 2048 + l * 64);
  uint32_t acc = 0;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int nb = 0; nb < 4; ++nb) {
      uint32_t r0, r1, r2, r3;
      asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(base + ((nb ^ ((l >> 1) & 3)) << 4)));
      acc += r0 ^ r1 ^ r2 ^ r3;
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (float)(t1 - t0);
  if (acc == 0x12345u) out[1] = 1;
}
int main() {
  float* out; cudaMalloc(&out, 16);
  float h;
  for (int warps : {1, 4, 8, 16}) for (int chains : {1, 4}) {
    const int iters = 2000;
    k_hmma<<<148, warps * 32>>>(out, iters, chains); cudaDeviceSynchronize();
    k_hmma<<<148, warps * 32>>>(out, iters, chains); cudaDeviceSynchronize();
    cudaMemcpy(&h, out, 4, cudaMemcpyDeviceToHost);
    const double per_smsp = (double)iters * chains * warps / 4.0;  // HMMA per SM sub-partition
    printf("HMMA m16n8k16 bf16: %2d warps/SM, %d independent chains: %.1f clk per HMMA per warp, %.1f clk per HMMA per SMSP -> %.0f FMA/clk/SM\n",
           warps, chains, h / (iters * chains), h / per_smsp, 2048.0 * 4 / (h / per_smsp));
  }
  for (int warps : {1, 8}) {
    const int iters = 2000;
    k_ldsm<<<148, warps * 32>>>(out, iters); cudaDeviceSynchronize();
    cudaMemcpy(&h, out, 4, cudaMemcpyDeviceToHost);
    printf("ldmatrix.x4.trans: %d warps/SM: %.1f clk per LDSM per warp\n", warps, h / (iters * 4));
  }
  return 0;
}
