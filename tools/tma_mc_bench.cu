// Microbenchmark: how many bytes/clk can one SM ingest from L2 through TMA, with and without
// cluster multicast?  (Design question behind profiles/r01_notes.md "W2 streaming".)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_mc_bench tools/tma_mc_bench.cu
// Every CTA consumes the same 4.4 MB bf16 matrix (4352 x 512) repeatedly as 16 KB stages
// (128 rows x 64 k, SWIZZLE_128B). CS = cluster size; with CS > 1 each CTA loads 1/CS of every
// stage (128/CS rows) and multicasts it to all CTAs of the cluster.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include "../rnnt-transducer_b200/csrc/sm100_ptx.cuh"
using namespace sm100;

constexpr int STAGES = 5;
constexpr int STAGE_BYTES = 128 * 128;

template <int CS>
__global__ void __launch_bounds__(128, 1)
bench_kernel(const __grid_constant__ CUtensorMap map, int iters, int rows_total) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  const uint32_t rank = (CS > 1) ? cluster_ctarank() : 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], CS); }
    fence_barrier_init();
  }
  __syncthreads();
  if (CS > 1) cluster_sync_all();
  const int n_vt = rows_total / 128;
  if (threadIdx.x == 0) {  // producer
    uint32_t stage = 0, phase = 0;
    for (int it = 0; it < iters; ++it) {
      const int vt = it % n_vt;
      for (int kb = 0; kb < 8; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full[stage], STAGE_BYTES);
        if (CS == 1) {
          tma_load_2d(smem + stage * STAGE_BYTES, &map, &full[stage], kb * 64, vt * 128);
        } else {
          constexpr int R = 128 / CS;
          const uint16_t mask = (uint16_t)((1u << CS) - 1);
          asm volatile(
              "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
              " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem + stage * STAGE_BYTES + rank * R * 128)),
              "l"(&map), "r"(smem_u32(&full[stage])), "r"(kb * 64), "r"(vt * 128 + (int)rank * R), "h"(mask)
              : "memory");
        }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (threadIdx.x == 32) {  // consumer: release the slot in every CTA of the cluster
    uint32_t stage = 0, phase = 0;
    for (int it = 0; it < iters * 8; ++it) {
      mbar_wait(&full[stage], phase);
      if (CS == 1) mbar_arrive(&empty[stage]);
      else
        for (int r = 0; r < CS; ++r) mbar_arrive_cluster(mapa_u32(smem_u32(&empty[stage]), r));
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  }
  __syncthreads();
  if (CS > 1) cluster_sync_all();
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int CS>
void run(EncodeTiledFn enc, void* w, int grid, int iters) {
  CUtensorMap map;
  cuuint64_t dims[2] = {512, 4352};
  cuuint64_t strides[1] = {1024};
  cuuint32_t box[2] = {64, (cuuint32_t)(128 / CS)};
  cuuint32_t es[2] = {1, 1};
  if (enc(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)) {
    printf("encode failed\n");
    return;
  }
  const int smem = STAGES * STAGE_BYTES + 256;
  cudaFuncSetAttribute(bench_kernel<CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid / CS * CS);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CS;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    cudaError_t e = cudaLaunchKernelEx(&cfg, bench_kernel<CS>, map, iters, 4352);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    if (e != cudaSuccess || cudaGetLastError() != cudaSuccess) { printf("launch failed CS=%d\n", CS); return; }
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double bytes_per_cta = (double)iters * 8 * STAGE_BYTES;
    printf("CS=%d grid=%d: %.3f ms, %.1f GB/s per SM received, %.2f TB/s received chip-wide, L2 reads %.2f TB/s\n", CS,
           grid / CS * CS, ms, bytes_per_cta / ms * 1e-6, bytes_per_cta * (grid / CS * CS) / ms * 1e-9,
           bytes_per_cta * (grid / CS * CS) / CS / ms * 1e-9);
  }
}

int main(int argc, char** argv) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaFree(0);
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  EncodeTiledFn enc = (EncodeTiledFn)fn;
  void* w;
  cudaMalloc(&w, 4352 * 512 * 2);
  cudaMemset(w, 0, 4352 * 512 * 2);
  const int iters = 2000;
  for (int grid : {148, 64, 16}) {
    run<1>(enc, w, grid, iters);
    run<2>(enc, w, grid, iters);
    run<4>(enc, w, grid, iters);
  }
  return 0;
}
