// bf16 GEMM on the 5th-generation tensor cores, hand-written for sm_100a:   C[M, N] = A[M, K] . W[N, K]^T (+ bias[N])
//
// The projection GEMMs of the SANA block (to_q | to_k | to_v on the fused [3C, C] weight, to_out, the two 1x1
// convolutions of GLUMBConv run token-major; diffusers Attention / GLUMBConv as used by
// utils/patch_sana_attention_layers.py:84-113) are ~65 % of the training step and were library calls (cuBLAS nvjet).
// This is the repo's own kernel for the `x @ W^T` form (torch F.linear: A = activations [tokens, in], W = weight
// [out, in], both K-major) - SURVEY.md §8(f) N1 / R2(b): the first step towards a projection whose epilogue is the
// linear-attention phase 1.
//
// Design (B200: 148 SMs = 74 CTA pairs, 256 KB TMEM and 227 KB shared memory per SM):
//   * cta_group::2: a cluster of two CTAs computes one 256 x BN output tile.  Each CTA stages its own 128 rows of A
//     and HALF of the W rows of the tile (BN / 2); one tcgen05.mma (M = 256, N = BN, K = 16) issued by the leader CTA
//     reads both halves, so every operand byte is staged once per pair.
//   * TMA (cp.async.bulk.tensor, SWIZZLE_128B) into a ring of kStages x {A 16 KB, W 16 KB} per CTA; both CTAs' loads
//     signal the LEADER's `full` barrier (mapa address), the MMA's tcgen05.commit frees the slot in both CTAs
//     (multicast arrive).
//   * accumulators in TMEM, two buffers of BN fp32 columns: the epilogue of tile i (tcgen05.ld -> +bias -> bf16 ->
//     256-bit global stores, four warps per CTA) overlaps the main loop of tile i + 1.
//   * persistent: grid = 74 clusters (or fewer tiles), tiles walked N-fastest so that concurrently running clusters
//     share A rows and the whole of W stays in the 126 MB L2.
// Warp roles per CTA (256 threads): 0 = TMA producer, 1 = tcgen05 issuer (leader CTA only), 2 = TMEM allocation,
// 4..7 = epilogue (TMEM lane quadrant = warp % 4).
// Edges: rows / columns / K beyond the matrix are zero-filled by TMA on load and predicated off on store.
#include "common.cuh"
#include "sm100_pair.cuh"

namespace yat {
namespace {
using namespace sm100;

constexpr int kBM = 128;                 // rows of A per CTA (256 per pair)
constexpr int kBK = 64;                  // K elements per stage = one 128-byte swizzle row
constexpr int kABytes = kBM * kBK * 2;   // 16 KB
constexpr int kWSlot = 128 * kBK * 2;    // 16 KB slot for up to 128 W rows (BN / 2)
constexpr int kStageBytes = kABytes + kWSlot;
constexpr int kMaxGemmStages = 7;
constexpr int kGemmThreads = 256;
constexpr int kAccStride = 256;          // TMEM columns between the two accumulators

template <int STAGES>
struct GemmSmem {
  uint8_t stage[STAGES][kStageBytes];            // 1024-aligned
  uint64_t full[STAGES], empty[STAGES];
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};
template <int STAGES>
constexpr int gemm_smem_bytes() { return (int)sizeof(GemmSmem<STAGES>) + 1024; }
static_assert(gemm_smem_bytes<kMaxGemmStages>() <= 232448, "GEMM shared memory exceeds the 227 KB per-CTA limit");

struct GemmMaps {
  CUtensorMap a, w;
};

// BN: tile width (multiple of 32, <= 256; BN / 2 W rows per CTA)
template <int BN, int kGemmStages>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tn_sm100_kernel(const __grid_constant__ GemmMaps maps, const __nv_bfloat16* __restrict__ bias,
                          __nv_bfloat16* __restrict__ c, int M, int N, int K, int64_t ldc) {
  static_assert(BN % 32 == 0 && BN <= 256 && BN >= 64, "tile width");
  constexpr int kHalfN = BN / 2;
  constexpr uint32_t kIdesc = make_idesc_bf16(256, BN, 0, 0);          // A, W both K-major
  constexpr uint32_t kTxBytes = 2u * (kABytes + kHalfN * kBK * 2);     // both CTAs' loads of one stage

  extern __shared__ uint8_t smem_raw[];
  using Smem = GemmSmem<kGemmStages>;
  Smem& S = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int n_clusters = gridDim.x >> 1, cluster_id = blockIdx.x >> 1;
  const int m_blocks = (M + 2 * kBM - 1) / (2 * kBM), n_blocks = (N + BN - 1) / BN;
  const int n_tiles = m_blocks * n_blocks;
  const int k_blocks = (K + kBK - 1) / kBK;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kGemmStages; ++s) {
      mbar_init(smem_u32(&S.full[s]), 1);            // the leader's producer arrives once per use (+ transaction bytes)
      mbar_init(smem_u32(&S.empty[s]), 1);           // tcgen05.commit (multicast)
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&S.tmem_full[a]), 1);       // tcgen05.commit (multicast)
      mbar_init(smem_u32(&S.tmem_empty[a]), 8);      // 4 epilogue warps x 2 CTAs (leader's copy is the one waited on)
    }
    fence_barrier_init();
    tma_prefetch_desc(&maps.a);
    tma_prefetch_desc(&maps.w);
  }
  if (warp == 2) {
    tmem_alloc2(smem_u32(&S.tmem_base), 512);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();                                // barriers initialised and TMEM allocated in both CTAs
  tc_fence_after();
  const uint32_t tmem = S.tmem_base;

  if (warp == 0) {
    // ======================= TMA producer (both CTAs) =======================
    if (lane == 0) {
      uint32_t it = 0;
      for (int tile = cluster_id; tile < n_tiles; tile += n_clusters) {
        const int m0 = (tile / n_blocks) * (2 * kBM) + (int)rank * kBM;
        const int n0 = (tile % n_blocks) * BN + (int)rank * kHalfN;
        for (int kb = 0; kb < k_blocks; ++kb, ++it) {
          const uint32_t s = it % kGemmStages, ph = (it / kGemmStages) & 1;
          mbar_wait(smem_u32(&S.empty[s]), ph ^ 1);
          const uint32_t full_leader = mapa_shared(smem_u32(&S.full[s]), 0);
          if (leader) mbar_arrive_expect_tx(smem_u32(&S.full[s]), kTxBytes);
          const uint32_t dst = smem_u32(&S.stage[s][0]);
          tma_load_2d_pair(dst, &maps.a, full_leader, kb * kBK, m0);
          tma_load_2d_pair(dst + kABytes, &maps.w, full_leader, kb * kBK, n0);
        }
      }
    }
  } else if (warp == 1) {
    // ======================= tcgen05 issuer (leader CTA, one thread) =======================
    if (leader && lane == 0) {
      uint32_t it = 0, local = 0;
      for (int tile = cluster_id; tile < n_tiles; tile += n_clusters, ++local) {
        const uint32_t acc = local & 1, use = local >> 1;
        mbar_wait(smem_u32(&S.tmem_empty[acc]), (use & 1) ^ 1);      // the epilogues of both CTAs have drained it
        tc_fence_after();
        const uint32_t d = tmem + acc * kAccStride;
        for (int kb = 0; kb < k_blocks; ++kb, ++it) {
          const uint32_t s = it % kGemmStages, ph = (it / kGemmStages) & 1;
          mbar_wait(smem_u32(&S.full[s]), ph);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(&S.stage[s][0]), w_addr = a_addr + kABytes;
#pragma unroll
          for (int ks = 0; ks < kBK / 16; ++ks) {
            // K-major SWIZZLE_128B: 8-row groups 1 KB apart; a K step of 16 elements = 32 B inside the 128 B row
            const uint64_t da = make_smem_desc(a_addr + ks * 32, 16, 1024, kLayoutSw128);
            const uint64_t db = make_smem_desc(w_addr + ks * 32, 16, 1024, kLayoutSw128);
            umma_bf16_pair(d, da, db, kIdesc, (kb | ks) != 0);
          }
          tc_commit_pair(smem_u32(&S.empty[s]));                     // slot free in both CTAs when these MMAs are done
        }
        tc_commit_pair(smem_u32(&S.tmem_full[acc]));                 // accumulator complete: wake both epilogues
      }
    }
  } else if (warp >= 4) {
    // ======================= epilogue (both CTAs): TMEM -> +bias -> bf16 -> global =======================
    const int quad = warp & 3;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const uint32_t empty_leader = mapa_shared(smem_u32(&S.tmem_empty[0]), 0);
    uint32_t local = 0;
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters, ++local) {
      const uint32_t acc = local & 1, use = local >> 1;
      const int row = (tile / n_blocks) * (2 * kBM) + (int)rank * kBM + quad * 32 + lane;
      const int n0 = (tile % n_blocks) * BN;
      mbar_wait(smem_u32(&S.tmem_full[acc]), use & 1);
      tc_fence_after();
      const uint32_t t_addr = tmem + lane_base + acc * kAccStride;
      __nv_bfloat16* crow = c + (int64_t)row * ldc + n0;
#pragma unroll 1
      for (int cc = 0; cc < BN; cc += 32) {
        uint32_t r[32];
        tmem_ld32(t_addr + cc, r);
        tmem_ld_wait();
        if (row < M && n0 + cc < N) {
          const int ncol = N - (n0 + cc);                             // valid columns from here (>= 1)
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            uint32_t w[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float lo = __uint_as_float(r[g * 16 + 2 * j]), hi = __uint_as_float(r[g * 16 + 2 * j + 1]);
              if (bias != nullptr) {
                const int col = n0 + cc + g * 16 + 2 * j;
                if (col + 1 < N) {
                  const uint32_t bw = __ldg(reinterpret_cast<const uint32_t*>(bias + col));
                  lo += bf16_lo(bw);
                  hi += bf16_hi(bw);
                } else if (col < N) {
                  lo += __bfloat162float(bias[col]);
                }
              }
              w[j] = pack_bf16(lo, hi);
            }
            const int c0 = g * 16;
            if (c0 + 16 <= ncol) {
              st_global_v8(crow + cc + c0, w);
            } else {
              for (int j = 0; j < 16 && c0 + j < ncol; ++j)
                crow[cc + c0 + j] = __ushort_as_bfloat16((unsigned short)((w[j >> 1] >> ((j & 1) * 16)) & 0xffffu));
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(empty_leader + acc * 8);     // uint64_t barriers are 8 bytes apart
    }
  }

  tc_fence_before();
  cluster_sync_all();                                // nobody exits (or frees TMEM) while the peer still needs this CTA
  if (warp == 2) tmem_dealloc2(tmem, 512);
}

}  // namespace

// host: 2-D bf16 tensor map {cols, rows} with row pitch `ld` elements, box {64, box_rows}, 128-byte swizzle
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_rows) {
  // a 3-D map with a unit batch dimension is a 2-D map the 2-D load instruction cannot address; encode rank 2 directly
  typedef CUresult (*PFN)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                          const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                          CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static PFN enc = nullptr;
  if (!enc) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      enc = (PFN)p;
  }
  YAT_REQUIRE(enc != nullptr, YAT_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {ld * 2};
  const cuuint32_t box[2] = {64, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = CUDA_ERROR_UNKNOWN;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT) break;
    cudaFree(nullptr);     // first driver-API call on this thread (autograd worker): bind the primary context, retry
  }
  YAT_REQUIRE(r == CUDA_SUCCESS, YAT_ERR_CUDA, "cuTensorMapEncodeTiled (2d) failed with CUresult %d", (int)r);
  return YAT_OK;
}

template <int BN, int STAGES>
static int launch_gemm(const GemmMaps& maps, const void* bias, void* c, int M, int N, int K, int64_t ldc,
                       cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    const cudaError_t e = cudaFuncSetAttribute(gemm_bf16_tn_sm100_kernel<BN, STAGES>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, gemm_smem_bytes<STAGES>());
    YAT_REQUIRE(e == cudaSuccess, YAT_ERR_CUDA, "yat_gemm_bf16_tn: cannot reserve %d B of shared memory (%s)",
                gemm_smem_bytes<STAGES>(), cudaGetErrorString(e));
    attr_set = true;
  }
  const int tiles = ((M + 255) / 256) * ((N + BN - 1) / BN);
  const int clusters = tiles < kNumSMs / 2 ? tiles : kNumSMs / 2;
  gemm_bf16_tn_sm100_kernel<BN, STAGES><<<2 * clusters, kGemmThreads, gemm_smem_bytes<STAGES>(), st>>>(
      maps, (const __nv_bfloat16*)bias, (__nv_bfloat16*)c, M, N, K, ldc);
  return check_launch("gemm_bf16_tn_sm100");
}

}  // namespace yat

extern "C" int yat_gemm_bf16_tn(const void* a, const void* w, const void* bias, void* c, int64_t M, int N, int K,
                                int64_t lda, int64_t ldw, int64_t ldc, int tile_n, void* stream) {
  using namespace yat;
  YAT_REQUIRE(a && w && c, YAT_ERR_BAD_ARG, "yat_gemm_bf16_tn: null pointer argument");
  YAT_REQUIRE(M > 0 && N > 0 && K > 0 && M < (1ll << 31), YAT_ERR_BAD_ARG, "yat_gemm_bf16_tn: bad dimensions");
  YAT_REQUIRE(K % 8 == 0 && lda % 8 == 0 && ldw % 8 == 0 && lda >= K && ldw >= K, YAT_ERR_BAD_ARG,
              "yat_gemm_bf16_tn: K, lda, ldw must be multiples of 8 elements (16-byte TMA rows), pitches >= K");
  YAT_REQUIRE(ldc >= N && ldc % 16 == 0, YAT_ERR_BAD_ARG,
              "yat_gemm_bf16_tn: ldc=%lld must be >= N and a multiple of 16 elements (256-bit stores)", (long long)ldc);
  YAT_REQUIRE((((uintptr_t)a | (uintptr_t)w) & 15u) == 0 && ((uintptr_t)c & 31u) == 0 &&
                  (bias == nullptr || ((uintptr_t)bias & 3u) == 0),
              YAT_ERR_BAD_ARG, "yat_gemm_bf16_tn: a / w must be 16-byte aligned, c 32-byte, bias 4-byte");
  int stages = 6;
  if (tile_n >= 1000) { stages = tile_n / 1000; tile_n %= 1000; }           // development: tile_n + 1000 * stages
  if (tile_n == 0) tile_n = 256;     // measured: 256 beats 224 even where 224 divides N (profiles/r02_gemm_ab.json)
  YAT_REQUIRE(stages >= 4 && stages <= 7, YAT_ERR_UNSUPPORTED, "yat_gemm_bf16_tn: %d stages not supported", stages);
  YAT_REQUIRE(tile_n == 256 || tile_n == 224 || tile_n == 128, YAT_ERR_UNSUPPORTED,
              "yat_gemm_bf16_tn: tile_n %d not supported (256, 224, 128)", tile_n);
  GemmMaps maps;
  int rc;
  if ((rc = make_tmap_2d_bf16(&maps.a, a, (uint64_t)K, (uint64_t)M, (uint64_t)lda, kBM)) != YAT_OK) return rc;
  if ((rc = make_tmap_2d_bf16(&maps.w, w, (uint64_t)K, (uint64_t)N, (uint64_t)ldw, (uint32_t)tile_n / 2)) != YAT_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
#define YAT_GEMM_CASE(BN, ST) if (tile_n == BN && stages == ST) return launch_gemm<BN, ST>(maps, bias, c, (int)M, N, K, ldc, st)
  YAT_GEMM_CASE(256, 6); YAT_GEMM_CASE(224, 6); YAT_GEMM_CASE(128, 6);
  YAT_GEMM_CASE(256, 7); YAT_GEMM_CASE(224, 7); YAT_GEMM_CASE(256, 5); YAT_GEMM_CASE(224, 5); YAT_GEMM_CASE(224, 4);
#undef YAT_GEMM_CASE
  YAT_REQUIRE(false, YAT_ERR_UNSUPPORTED, "yat_gemm_bf16_tn: tile_n %d with %d stages is not built", tile_n, stages);
  return YAT_ERR_UNSUPPORTED;
}
