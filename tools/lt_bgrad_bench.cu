// Experiment: can cublasLt produce d b2 (column sums of dS) as a BGRADA epilogue of the dW2 GEMM?
//   nvcc -O3 -o tools/lt_bgrad_bench tools/lt_bgrad_bench.cu -lcublasLt
// D (col-major Vds x H, fp32) = A (col-major Vds x rows  == row-major dS[rows][Vds], bf16)
//                             . B^T (B col-major H x rows == row-major h[rows][H], bf16)
// BGRADA: bias[m] = sum_k A[m][k] = column sums of dS.
#include <cublasLt.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include <math.h>

#define CK(x) do { auto e_ = (x); if (e_ != 0) { printf("FAIL %s = %d line %d\n", #x, (int)e_, __LINE__); exit(1); } } while (0)

__global__ void fill(__nv_bfloat16* p, size_t n, unsigned seed, float scale) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    unsigned x = (unsigned)i * 2654435761u + seed;
    x ^= x >> 15; x *= 2246822519u; x ^= x >> 13;
    p[i] = __float2bfloat16(((x & 0xffff) / 65536.f - 0.5f) * scale);
  }
}

struct Run { float ms; int ok; };

Run run(cublasLtHandle_t lt, int rows, int Vds, int H, int epi, float beta, void* ws, size_t ws_bytes,
        __nv_bfloat16* ds, __nv_bfloat16* h, float* d, float* bias, int iters) {
  cublasLtMatmulDesc_t op;
  CK(cublasLtMatmulDescCreate(&op, CUBLAS_COMPUTE_32F, CUDA_R_32F));
  cublasOperation_t tn = CUBLAS_OP_N, tt = CUBLAS_OP_T;
  CK(cublasLtMatmulDescSetAttribute(op, CUBLASLT_MATMUL_DESC_TRANSA, &tn, sizeof(tn)));
  CK(cublasLtMatmulDescSetAttribute(op, CUBLASLT_MATMUL_DESC_TRANSB, &tt, sizeof(tt)));
  if (epi) {
    cublasLtEpilogue_t e = CUBLASLT_EPILOGUE_BGRADA;
    CK(cublasLtMatmulDescSetAttribute(op, CUBLASLT_MATMUL_DESC_EPILOGUE, &e, sizeof(e)));
    CK(cublasLtMatmulDescSetAttribute(op, CUBLASLT_MATMUL_DESC_BIAS_POINTER, &bias, sizeof(bias)));
    cudaDataType_t bt = CUDA_R_32F;
    CK(cublasLtMatmulDescSetAttribute(op, CUBLASLT_MATMUL_DESC_BIAS_DATA_TYPE, &bt, sizeof(bt)));
  }
  cublasLtMatrixLayout_t la, lb, ld;
  CK(cublasLtMatrixLayoutCreate(&la, CUDA_R_16BF, Vds, rows, Vds));   // A: Vds x rows, ld Vds
  CK(cublasLtMatrixLayoutCreate(&lb, CUDA_R_16BF, H, rows, H));       // B: H x rows, ld H (op T -> rows x H)
  CK(cublasLtMatrixLayoutCreate(&ld, CUDA_R_32F, Vds, H, Vds));       // D: Vds x H, ld Vds
  cublasLtMatmulPreference_t pref;
  CK(cublasLtMatmulPreferenceCreate(&pref));
  CK(cublasLtMatmulPreferenceSetAttribute(pref, CUBLASLT_MATMUL_PREF_MAX_WORKSPACE_BYTES, &ws_bytes, sizeof(ws_bytes)));
  cublasLtMatmulHeuristicResult_t res[8];
  int n = 0;
  cublasStatus_t st = cublasLtMatmulAlgoGetHeuristic(lt, op, la, lb, ld, ld, pref, 8, res, &n);
  Run r = {0.f, 0};
  if (st != CUBLAS_STATUS_SUCCESS || n == 0) { printf("  no algorithm (status %d, n %d)\n", (int)st, n); return r; }
  float alpha = 1.f;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int a = 0; a < n && a < 4; ++a) {
    st = cublasLtMatmul(lt, op, &alpha, ds, la, h, lb, &beta, d, ld, d, ld, &res[a].algo, ws, ws_bytes, 0);
    if (st != CUBLAS_STATUS_SUCCESS) { printf("  algo %d failed %d\n", a, (int)st); continue; }
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int i = 0; i < iters; ++i)
      cublasLtMatmul(lt, op, &alpha, ds, la, h, lb, &beta, d, ld, d, ld, &res[a].algo, ws, ws_bytes, 0);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    ms /= iters;
    printf("  algo %d: %.3f ms (%.0f TFLOP/s) ws %zu waves %.2f\n", a, ms, 2.0 * rows * Vds * H / ms * 1e-9,
           res[a].workspaceSize, res[a].wavesCount);
    if (ms < best) best = ms;
    r.ok = 1;
  }
  r.ms = best;
  return r;
}

int main() {
  cublasLtHandle_t lt;
  CK(cublasLtCreate(&lt));
  const int Vds = 4288, H = 512;
  size_t ws_bytes = 256u << 20;
  void* ws;
  CK(cudaMalloc(&ws, ws_bytes));
  // correctness at a small size
  {
    const int rows = 1024;
    __nv_bfloat16 *ds, *h; float *d, *bias;
    CK(cudaMalloc(&ds, (size_t)rows * Vds * 2)); CK(cudaMalloc(&h, (size_t)rows * H * 2));
    CK(cudaMalloc(&d, (size_t)Vds * H * 4)); CK(cudaMalloc(&bias, Vds * 4));
    fill<<<1024, 256>>>(ds, (size_t)rows * Vds, 1u, 1.f);
    fill<<<1024, 256>>>(h, (size_t)rows * H, 2u, 1.f);
    cudaMemset(bias, 0, Vds * 4);
    printf("small rows=%d BGRADA:\n", rows);
    Run r = run(lt, rows, Vds, H, 1, 0.f, ws, ws_bytes, ds, h, d, bias, 1);
    if (r.ok) {
      std::vector<__nv_bfloat16> hds((size_t)rows * Vds);
      std::vector<float> hb(Vds);
      cudaMemcpy(hds.data(), ds, hds.size() * 2, cudaMemcpyDeviceToHost);
      cudaMemcpy(hb.data(), bias, Vds * 4, cudaMemcpyDeviceToHost);
      double maxerr = 0;
      for (int v = 0; v < Vds; v += 97) {
        double s = 0;
        for (int r2 = 0; r2 < rows; ++r2) s += __bfloat162float(hds[(size_t)r2 * Vds + v]);
        maxerr = fmax(maxerr, fabs(s - hb[v]));
      }
      printf("  bias-gradient max abs err vs column sums: %.3e\n", maxerr);
    }
    cudaFree(ds); cudaFree(h); cudaFree(d); cudaFree(bias);
  }
  // timing at one dS chunk of the AISHELL shape
  {
    const int rows = 3157 * 128;
    __nv_bfloat16 *ds, *h; float *d, *bias;
    CK(cudaMalloc(&ds, (size_t)rows * Vds * 2)); CK(cudaMalloc(&h, (size_t)rows * H * 2));
    CK(cudaMalloc(&d, (size_t)Vds * H * 4)); CK(cudaMalloc(&bias, Vds * 4));
    fill<<<4096, 256>>>(ds, (size_t)rows * Vds, 1u, 0.01f);
    fill<<<4096, 256>>>(h, (size_t)rows * H, 2u, 1.f);
    cudaMemset(d, 0, (size_t)Vds * H * 4);
    printf("rows=%d plain, beta=0:\n", rows);
    run(lt, rows, Vds, H, 0, 0.f, ws, ws_bytes, ds, h, d, bias, 5);
    printf("rows=%d plain, beta=1:\n", rows);
    run(lt, rows, Vds, H, 0, 1.f, ws, ws_bytes, ds, h, d, bias, 5);
    printf("rows=%d BGRADA, beta=0:\n", rows);
    run(lt, rows, Vds, H, 1, 0.f, ws, ws_bytes, ds, h, d, bias, 5);
    printf("rows=%d BGRADA, beta=1:\n", rows);
    run(lt, rows, Vds, H, 1, 1.f, ws, ws_bytes, ds, h, d, bias, 5);
  }
  return 0;
}
