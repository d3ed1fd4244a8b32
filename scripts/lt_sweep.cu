// Experiment (not part of the product library): sustained, power-capped throughput of every cuBLASLt algorithm the
// heuristic returns for the GEMM shapes that dominate the SANA-1.6B step (profiles/r02e_launches.md), to see whether the
// library's first choice is also the most frugal one at the 1 kW cap.  Column-major shapes as torch hands them to cuBLAS
// (the TunableOp signature: <transa><transb>_m_n_k_ld_lda_ldb_ldc).
//   nvcc -O2 -std=c++17 -o gpurun_out/lt_sweep scripts/lt_sweep.cu -lcublasLt && gpurun_out/lt_sweep [seconds_per_algo]
#include <cublasLt.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                      \
  do {                                                                             \
    auto e_ = (x);                                                                 \
    if ((int)e_ != 0) { printf("error %d at %s:%d\n", (int)e_, __FILE__, __LINE__); exit(1); } \
  } while (0)

// operands with realistic bit entropy (board power, and with it the capped clock, depends on the data): uniform in [-1, 1)
__global__ void fill_uniform(__nv_bfloat16* p, size_t n, uint32_t seed) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    uint32_t h = (uint32_t)i * 2654435761u ^ seed;
    h ^= h >> 16; h *= 0x7feb352du; h ^= h >> 15; h *= 0x846ca68bu; h ^= h >> 16;
    p[i] = __float2bfloat16((h >> 8) * (1.0f / 8388608.0f) - 1.0f);
  }
}

struct Shape { const char* name; char ta, tb; int64_t m, n, k, lda, ldb, ldc; };

int main(int argc, char** argv) {
  const double seconds = argc > 1 ? atof(argv[1]) : 1.0;
  const Shape shapes[] = {
      {"qkv fwd", 'T', 'N', 6720, 16384, 2240, 2240, 2240, 6720},
      {"qkv dgrad", 'N', 'N', 2240, 16384, 6720, 2240, 6720, 2240},
      {"qkv wgrad", 'N', 'T', 2240, 6720, 16384, 2240, 6720, 2240},
      {"conv_inverted fwd", 'T', 'N', 11200, 16384, 2240, 2240, 2240, 11200},
      {"conv_inverted dgrad", 'N', 'N', 2240, 16384, 11200, 2240, 11200, 2240},
      {"conv_inverted wgrad", 'N', 'T', 2240, 11200, 16384, 2240, 11200, 2240},
      {"conv_point fwd", 'T', 'N', 2240, 16384, 5600, 5600, 5600, 2240},
      {"conv_point dgrad", 'N', 'N', 5600, 16384, 2240, 5600, 2240, 5600},
      {"conv_point wgrad", 'N', 'T', 5600, 2240, 16384, 5600, 2240, 5600},
      {"to_out fwd", 'T', 'N', 2240, 16384, 2240, 2240, 2240, 2240},
      {"to_out dgrad", 'N', 'N', 2240, 16384, 2240, 2240, 2240, 2240},
      {"to_out wgrad", 'N', 'T', 2240, 2240, 16384, 2240, 2240, 2240},
  };
  cublasLtHandle_t lt;
  CK(cublasLtCreate(&lt));
  const size_t ws_bytes = 64u << 20;
  void* ws;
  CK(cudaMalloc(&ws, ws_bytes));
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  const float alpha = 1.f, beta = 0.f;
  for (const Shape& s : shapes) {
    const int64_t a_rows = s.ta == 'N' ? s.m : s.k, a_cols = s.ta == 'N' ? s.k : s.m;
    const int64_t b_rows = s.tb == 'N' ? s.k : s.n, b_cols = s.tb == 'N' ? s.n : s.k;
    __nv_bfloat16 *A, *B, *C;
    CK(cudaMalloc(&A, s.lda * a_cols * 2));
    CK(cudaMalloc(&B, s.ldb * b_cols * 2));
    CK(cudaMalloc(&C, s.ldc * s.n * 2));
    fill_uniform<<<1184, 256>>>(A, (size_t)(s.lda * a_cols), 1u);
    fill_uniform<<<1184, 256>>>(B, (size_t)(s.ldb * b_cols), 2u);
    CK(cudaDeviceSynchronize());
    cublasLtMatmulDesc_t op;
    CK(cublasLtMatmulDescCreate(&op, CUBLAS_COMPUTE_32F, CUDA_R_32F));
    cublasOperation_t ta = s.ta == 'T' ? CUBLAS_OP_T : CUBLAS_OP_N, tb = s.tb == 'T' ? CUBLAS_OP_T : CUBLAS_OP_N;
    CK(cublasLtMatmulDescSetAttribute(op, CUBLASLT_MATMUL_DESC_TRANSA, &ta, sizeof(ta)));
    CK(cublasLtMatmulDescSetAttribute(op, CUBLASLT_MATMUL_DESC_TRANSB, &tb, sizeof(tb)));
    cublasLtMatrixLayout_t la, lb, lc;
    CK(cublasLtMatrixLayoutCreate(&la, CUDA_R_16BF, a_rows, a_cols, s.lda));
    CK(cublasLtMatrixLayoutCreate(&lb, CUDA_R_16BF, b_rows, b_cols, s.ldb));
    CK(cublasLtMatrixLayoutCreate(&lc, CUDA_R_16BF, s.m, s.n, s.ldc));
    cublasLtMatmulPreference_t pref;
    CK(cublasLtMatmulPreferenceCreate(&pref));
    CK(cublasLtMatmulPreferenceSetAttribute(pref, CUBLASLT_MATMUL_PREF_MAX_WORKSPACE_BYTES, &ws_bytes, sizeof(ws_bytes)));
    cublasLtMatmulHeuristicResult_t res[16];
    int found = 0;
    CK(cublasLtMatmulAlgoGetHeuristic(lt, op, la, lb, lc, lc, pref, 16, res, &found));
    const double flop = 2.0 * s.m * s.n * s.k;
    printf("== %s %c%c m=%lld n=%lld k=%lld: %d algorithms\n", s.name, s.ta, s.tb, (long long)s.m, (long long)s.n, (long long)s.k, found);
    double first = 0;
    if (found > 0) {   // bring the board to its power-capped state before the first measurement
      const auto t0 = std::chrono::steady_clock::now();
      while (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() < 2.0) {
        for (int j = 0; j < 50; ++j)
          cublasLtMatmul(lt, op, &alpha, A, la, B, lb, &beta, C, lc, C, lc, &res[0].algo, ws, res[0].workspaceSize, st);
        CK(cudaStreamSynchronize(st));
      }
    }
    for (int i = 0; i < found; ++i) {
      if (res[i].state != CUBLAS_STATUS_SUCCESS) continue;
      int id = -1, tile = -1, stages = -1, cluster = -1, splitk = -1;
      size_t w = 0;
      cublasLtMatmulAlgoConfigGetAttribute(&res[i].algo, CUBLASLT_ALGO_CONFIG_ID, &id, sizeof(id), &w);
      cublasLtMatmulAlgoConfigGetAttribute(&res[i].algo, CUBLASLT_ALGO_CONFIG_TILE_ID, &tile, sizeof(tile), &w);
      cublasLtMatmulAlgoConfigGetAttribute(&res[i].algo, CUBLASLT_ALGO_CONFIG_STAGES_ID, &stages, sizeof(stages), &w);
      cublasLtMatmulAlgoConfigGetAttribute(&res[i].algo, CUBLASLT_ALGO_CONFIG_CLUSTER_SHAPE_ID, &cluster, sizeof(cluster), &w);
      cublasLtMatmulAlgoConfigGetAttribute(&res[i].algo, CUBLASLT_ALGO_CONFIG_SPLITK_NUM, &splitk, sizeof(splitk), &w);
      auto run = [&]() {
        return cublasLtMatmul(lt, op, &alpha, A, la, B, lb, &beta, C, lc, C, lc, &res[i].algo, ws, res[i].workspaceSize, st);
      };
      if (run() != CUBLAS_STATUS_SUCCESS) { printf("   #%d launch failed\n", i); continue; }
      for (int j = 0; j < 20; ++j) run();
      CK(cudaStreamSynchronize(st));
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      long n = 0;
      const auto t0 = std::chrono::steady_clock::now();
      cudaEventRecord(e0, st);
      while (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() < seconds) {
        for (int j = 0; j < 50; ++j) run();
        n += 50;
        CK(cudaStreamSynchronize(st));
      }
      cudaEventRecord(e1, st);
      CK(cudaStreamSynchronize(st));
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double us = ms * 1e3 / n, tf = flop / (us * 1e-6) / 1e12;
      if (i == 0) first = us;
      printf("   #%-2d id=%-3d tile=%-4d stages=%-3d cluster=%-3d splitk=%-2d ws=%-9zu waves=%.2f  %8.1f us  %7.1f TFLOP/s  x%.3f of #0\n", i, id,
             tile, stages, cluster, splitk, res[i].workspaceSize, res[i].wavesCount, us, tf, first / us);
      fflush(stdout);
      cudaEventDestroy(e0); cudaEventDestroy(e1);
    }
    cublasLtMatmulPreferenceDestroy(pref);
    cublasLtMatrixLayoutDestroy(la); cublasLtMatrixLayoutDestroy(lb); cublasLtMatrixLayoutDestroy(lc);
    cublasLtMatmulDescDestroy(op);
    cudaFree(A); cudaFree(B); cudaFree(C);
  }
  return 0;
}
