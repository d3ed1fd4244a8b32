// to_q | to_k | to_v projection of SANA's linear attention with the attention's phase 1 as the GEMM epilogue
// (SURVEY.md §8(f) N1 / §7.3 R2(b)): one tcgen05 kernel computes
//
//   qkv[b, n, :] = x[b, n, :] . Wqkv^T                      [B, N, 3C] bf16  (q | k | v, what the backward reads)
//   KV_h = relu(k_h)^T v_h   (32 x 32)     Ks_h = sum_n relu(k_h)[n]         per (batch, head), as per-128-token partials
//
// so that the attention forward (yat_linattn_fwd_state, linattn_sm100.cu) only streams q and the 33 x 32 state: k and v
// never come back from HBM in the forward.  Reference: diffusers Attention.to_q/to_k/to_v + SanaLinearAttnProcessor2_0,
// reached from utils/patch_sana_attention_layers.py:94; with the projection inside, the attention is a dense
// contraction (493 GFLOP per SANA-1.6B layer) and sits on the tensor roofline instead of the HBM one.
//
// GEMM part = gemm_sm100.cu (cta_group::2 pairs, 256 x 192 tiles, TMA ring, two TMEM accumulators).  The pairing is
// what makes the fusion free of any weight re-layout: each CTA of a pair stages HALF of the tile's W rows, so for a
// k/v tile CTA 0 loads the 96 to_k rows of a 3-head group and CTA 1 the 96 to_v rows of the SAME heads - the tile
// is [k_g | v_g] for the pair's 256 tokens although the two row blocks are C rows apart in Wqkv.
//
// Epilogue of a k/v tile, per CTA (its 128 tokens): TMEM -> bf16 -> global (k, v as a plain GEMM would), and the same
// registers -> shared memory as tcgen05 operands: relu(k) as an MN-major A tile [128 tokens x 96(+32)], v as two
// MN-major B tiles [128 tokens x 64] (heads 0,1 | head 2, a ones column, zeros).  A second issuer thread runs, per
// chunk, one pair MMA over the token axis  D[256 x 128] = relu(K)^T [V0 | V1]:  rows = k columns of CTA r, columns
// [64 r, 64 r + 64) = CTA r's v chunk (the cross blocks pair tokens of different CTAs and are discarded, like the
// off-diagonal head blocks).  The ones column yields Ks.  The 32 x 32 diagonal blocks leave as fp32 partials
// part[b][128-token tile][head][33][32] (row 32 = Ks), summed in fixed order by the consumer: deterministic, no atomics.
// TMEM: accumulators [0,192) [192,384), KV chunk [384,512).
#include "common.cuh"
#include "sm100_pair.cuh"

namespace yat {
namespace {
using namespace sm100;

constexpr int kBM = 128;                       // tokens per CTA (256 per pair)
constexpr int kBK = 64;                        // K elements per stage = one 128-byte swizzle row
constexpr int kBN = 192;                       // tile width: 6 heads of q, or 3 heads of k + the same 3 heads of v
constexpr int kHalfN = kBN / 2;                // W rows staged per CTA
constexpr int kGroupHeads = 3;
constexpr int kABytes = kBM * kBK * 2;         // 16 KB
constexpr int kWBytes = kHalfN * kBK * 2;      // 12 KB
constexpr int kStageBytes = kABytes + kWBytes; // 28 KB (multiple of 1024: SWIZZLE_128B alignment)
constexpr int kStages = 5;
constexpr int kThreads = 256;
constexpr int kHalfBytes = kBM * 128;          // one [128 tokens][64 cols] operand tile
constexpr uint32_t kAccStride = kBN;           // TMEM columns between the two accumulators
constexpr uint32_t kColKV = 2 * kBN;           // 384: the [128 x 128] KV chunk
constexpr int kStateRow = 33 * 32;             // floats per (tile, head) partial: KV rows 0..31, Ks row 32

struct QkvSmem {
  uint8_t stage[kStages][kStageBytes];         // 1024-aligned
  uint8_t ka[2][kHalfBytes];                   // relu(k): columns 0..63 | 64..95 (64..127 of the second half are never read back)
  uint8_t vb[2][kHalfBytes];                   // chunk 0: v columns 0..63;  chunk 1: v columns 64..95, ones, zeros
  uint64_t full[kStages], empty[kStages];
  uint64_t tmem_full[2], tmem_empty[2];
  uint64_t stage_ready, kv_full, kv_empty;
  uint32_t tmem_base;
};
constexpr int kQkvSmemBytes = (int)sizeof(QkvSmem) + 1024;
static_assert(kQkvSmemBytes <= 232448, "qkv kernel exceeds the 227 KB shared-memory limit");
static_assert(kStageBytes % 1024 == 0, "stage alignment");

struct QkvMaps {
  CUtensorMap x, w;
};
struct QkvArgs {
  __nv_bfloat16* qkv;   // [B, N, 3C]
  float* part;          // [B, 2 * m_blocks, H, 33, 32]
  int B, N, H, C, Cin;
};

struct TileInfo {
  int b, mb, nb, g;
  bool kv;
};

__device__ __forceinline__ void st_global_v8(void* p, const uint32_t (&w)[8]) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]),
               "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
               : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
qkv_linattn_sm100_kernel(const __grid_constant__ QkvMaps maps, const QkvArgs a) {
  constexpr uint32_t kIdesc = make_idesc_bf16(256, kBN, 0, 0);          // x, W both K-major
  constexpr uint32_t kIdescKV = make_idesc_bf16(256, 128, 1, 1);        // relu(K)^T V over tokens: A, B MN-major
  constexpr uint32_t kTxBytes = 2u * kStageBytes;                       // both CTAs' loads of one stage

  extern __shared__ uint8_t smem_raw[];
  QkvSmem& S = *reinterpret_cast<QkvSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int n_clusters = gridDim.x >> 1, cluster_id = blockIdx.x >> 1;
  const int m_blocks = (a.N + 2 * kBM - 1) / (2 * kBM);
  const int nq = (a.C + kBN - 1) / kBN, groups = (a.H + kGroupHeads - 1) / kGroupHeads;
  const int n_blocks = nq + groups;
  const int n_tiles = a.B * m_blocks * n_blocks;
  const int k_blocks = (a.Cin + kBK - 1) / kBK;

  auto decode = [&](int tile) {
    TileInfo t;
    t.nb = tile % n_blocks;
    const int rest = tile / n_blocks;
    t.mb = rest % m_blocks;
    t.b = rest / m_blocks;
    t.kv = t.nb >= nq;
    t.g = t.nb - nq;
    return t;
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(smem_u32(&S.full[s]), 1);            // the leader's producer arrives once per use (+ transaction bytes)
      mbar_init(smem_u32(&S.empty[s]), 1);           // tcgen05.commit (multicast)
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&S.tmem_full[i]), 1);       // tcgen05.commit (multicast)
      mbar_init(smem_u32(&S.tmem_empty[i]), 8);      // 4 epilogue warps x 2 CTAs (the leader's copy is the one waited on)
    }
    mbar_init(smem_u32(&S.stage_ready), 8);          // operand tiles written in both CTAs (leader's copy)
    mbar_init(smem_u32(&S.kv_full), 1);              // tcgen05.commit (multicast): a KV chunk is in TMEM
    mbar_init(smem_u32(&S.kv_empty), 8);             // both CTAs have read the chunk (leader's copy)
    fence_barrier_init();
    tma_prefetch_desc(&maps.x);
    tma_prefetch_desc(&maps.w);
  }
  // constant part of the second V chunk: column 32 = 1 (-> Ks), columns 33..63 = 0; columns 0..31 are rewritten per tile
  for (int i = threadIdx.x; i < kBM * 4; i += kThreads) {
    const int r = i >> 2, c16 = 4 + (i & 3);
    const uint4 v = (i & 3) == 0 ? make_uint4(0x00003f80u, 0, 0, 0) : make_uint4(0, 0, 0, 0);   // bf16 1.0 in the low half
    *reinterpret_cast<uint4*>(&S.vb[1][r * 128 + ((c16 ^ (r & 7)) << 4)]) = v;
  }
  fence_proxy_async();
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
        const TileInfo t = decode(tile);
        const int tok0 = t.mb * (2 * kBM) + (int)rank * kBM;
        // W rows of this CTA's half: q tile -> consecutive rows; k/v tile -> to_k rows (rank 0) / to_v rows (rank 1) of group g
        const int wrow = t.kv ? (int)(rank + 1) * a.C + t.g * kHalfN : t.nb * kBN + (int)rank * kHalfN;
        for (int kb = 0; kb < k_blocks; ++kb, ++it) {
          const uint32_t s = it % kStages, ph = (it / kStages) & 1;
          mbar_wait(smem_u32(&S.empty[s]), ph ^ 1);
          const uint32_t full_leader = mapa_shared(smem_u32(&S.full[s]), 0);
          if (leader) mbar_arrive_expect_tx(smem_u32(&S.full[s]), kTxBytes);
          const uint32_t dst = smem_u32(&S.stage[s][0]);
          tma_load_3d_pair(dst, &maps.x, full_leader, kb * kBK, tok0, t.b);
          tma_load_2d_pair(dst + kABytes, &maps.w, full_leader, kb * kBK, wrow);
        }
      }
    }
  } else if (warp == 1) {
    // ======================= projection issuer (leader CTA, one thread) =======================
    if (leader && lane == 0) {
      uint32_t it = 0, local = 0;
      for (int tile = cluster_id; tile < n_tiles; tile += n_clusters, ++local) {
        const uint32_t acc = local & 1, use = local >> 1;
        mbar_wait(smem_u32(&S.tmem_empty[acc]), (use & 1) ^ 1);      // the epilogues of both CTAs have drained it
        tc_fence_after();
        const uint32_t d = tmem + acc * kAccStride;
        for (int kb = 0; kb < k_blocks; ++kb, ++it) {
          const uint32_t s = it % kStages, ph = (it / kStages) & 1;
          mbar_wait(smem_u32(&S.full[s]), ph);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(&S.stage[s][0]), w_addr = a_addr + kABytes;
#pragma unroll
          for (int ks = 0; ks < kBK / 16; ++ks) {
            const uint64_t da = make_smem_desc(a_addr + ks * 32, 16, 1024, kLayoutSw128);
            const uint64_t db = make_smem_desc(w_addr + ks * 32, 16, 1024, kLayoutSw128);
            umma_bf16_pair(d, da, db, kIdesc, (kb | ks) != 0);
          }
          tc_commit_pair(smem_u32(&S.empty[s]));                     // slot free in both CTAs when these MMAs are done
        }
        tc_commit_pair(smem_u32(&S.tmem_full[acc]));                 // accumulator complete: wake both epilogues
      }
    }
  } else if (warp == 3) {
    // ======================= KV issuer (leader CTA, one thread): D = relu(K)^T [V chunk] over the 128 tokens =======================
    if (leader && lane == 0) {
      uint32_t n_kv = 0;
      const uint32_t ka = smem_u32(&S.ka[0][0]);
      for (int tile = cluster_id; tile < n_tiles; tile += n_clusters) {
        if (!decode(tile).kv) continue;
        mbar_wait_cluster(smem_u32(&S.stage_ready), n_kv & 1);    // operand tiles of BOTH CTAs (the peer's via release.cluster)
        tc_fence_after();
#pragma unroll
        for (int chunk = 0; chunk < 2; ++chunk) {
          const uint32_t vb = smem_u32(&S.vb[chunk][0]);
#pragma unroll
          for (int ks = 0; ks < kBM / 16; ++ks) {
            // MN-major SWIZZLE_128B tiles [tokens][64 cols]: LBO = stride between 64-column halves, SBO = 8-token group
            const uint64_t da = make_smem_desc(ka + ks * 2048, kHalfBytes, 1024, kLayoutSw128);
            const uint64_t db = make_smem_desc(vb + ks * 2048, kHalfBytes, 1024, kLayoutSw128);
            umma_bf16_pair(tmem + kColKV, da, db, kIdescKV, ks != 0);
          }
          tc_commit_pair(smem_u32(&S.kv_full));
          mbar_wait(smem_u32(&S.kv_empty), (2 * n_kv + chunk) & 1);  // both CTAs have read the chunk out of TMEM
          tc_fence_after();
        }
        ++n_kv;
      }
    }
  } else if (warp >= 4) {
    // ======================= epilogue (both CTAs) =======================
    const int quad = warp & 3;
    const int row = quad * 32 + lane;                                // token row of the CTA's tile = TMEM lane
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const uint32_t empty_leader = mapa_shared(smem_u32(&S.tmem_empty[0]), 0);
    const uint32_t ready_leader = mapa_shared(smem_u32(&S.stage_ready), 0);
    const uint32_t kvempty_leader = mapa_shared(smem_u32(&S.kv_empty), 0);
    const int64_t ld = 3 * (int64_t)a.C;
    const int swz = row & 7;
    uint32_t local = 0, kv_ph = 0;
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters, ++local) {
      const TileInfo t = decode(tile);
      const uint32_t acc = local & 1, use = local >> 1;
      const int tok = t.mb * (2 * kBM) + (int)rank * kBM + row;
      const bool row_ok = tok < a.N;
      mbar_wait(smem_u32(&S.tmem_full[acc]), use & 1);
      tc_fence_after();
      const uint32_t t_addr = tmem + lane_base + acc * kAccStride;
      __nv_bfloat16* orow = a.qkv + ((int64_t)t.b * a.N + tok) * ld;
#pragma unroll 1
      for (int cc = 0; cc < kBN / 32; ++cc) {
        uint32_t r[32];
        tmem_ld32(t_addr + cc * 32, r);
        tmem_ld_wait();
        // global column of this 32-column chunk (one head wide; C is a multiple of 32)
        int col, inner;                                              // inner: column inside its q / k / v third
        if (!t.kv) { inner = t.nb * kBN + cc * 32; col = inner; }
        else       { inner = t.g * kHalfN + (cc % 3) * 32; col = (cc < 3 ? 1 : 2) * a.C + inner; }
        const bool col_ok = inner < a.C;
        uint32_t w[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) w[j] = pack_bf16(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
        if (row_ok && col_ok) {
          uint32_t w0[8], w1[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) { w0[j] = w[j]; w1[j] = w[8 + j]; }
          st_global_v8(orow + col, w0);
          st_global_v8(orow + col + 16, w1);
        }
        if (t.kv) {
          // the same values as tensor-core operands (rows beyond N contribute nothing: relu(k) = 0 there)
          const bool is_k = cc < 3;
          const int hc = cc % 3;                                     // head inside the group
          uint8_t* base = is_k ? &S.ka[hc >> 1][0] : &S.vb[hc >> 1][0];
          const int c16 = (hc & 1) * 4;
          const __nv_bfloat162 z = __float2bfloat162_rn(0.f);
#pragma unroll
          for (int g4 = 0; g4 < 4; ++g4) {
            uint4 v = make_uint4(w[4 * g4], w[4 * g4 + 1], w[4 * g4 + 2], w[4 * g4 + 3]);
            if (is_k) {
              __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
              for (int i = 0; i < 4; ++i) h[i] = __hmax2_nan(h[i], z);
            }
            if (!row_ok || !col_ok) v = make_uint4(0, 0, 0, 0);
            *reinterpret_cast<uint4*>(base + row * 128 + (((c16 + g4) ^ swz) << 4)) = v;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(empty_leader + acc * 8);     // accumulator drained (uint64_t barriers are 8 bytes apart)
      if (t.kv) {
        fence_proxy_async();                                         // operand tiles -> visible to the tensor core
        __syncwarp();
        if (lane == 0) mbar_arrive_remote(ready_leader);
        // This warp's TMEM lanes are the k columns of head `quad` of the group (quad 3: padding rows, nothing to do).
        const int head = t.g * kGroupHeads + quad;
        const bool head_ok = quad < kGroupHeads && head < a.H;
        float* prow = a.part + ((((int64_t)t.b * (2 * m_blocks) + 2 * t.mb + (int)rank) * a.H + head) * kStateRow);
        const uint32_t kv_addr = tmem + lane_base + kColKV + rank * 64;   // this CTA's v chunk columns
#pragma unroll 1
        for (int chunk = 0; chunk < 2; ++chunk) {
          mbar_wait(smem_u32(&S.kv_full), kv_ph & 1);
          ++kv_ph;
          tc_fence_after();
          // chunk 0 holds v heads 0, 1 (32 columns each); chunk 1 head 2 in columns 0..31 and Ks in column 32
          const bool mine = chunk == 0 ? quad < 2 : quad == 2;
          if (mine && head_ok) {
            uint32_t r[32];
            tmem_ld32(kv_addr + (chunk == 0 ? quad * 32 : 0), r);
            tmem_ld_wait();
            float4* dst = reinterpret_cast<float4*>(prow + lane * 32);   // KV_head[dk = lane][dv = 0..31]
#pragma unroll
            for (int j = 0; j < 8; ++j)
              dst[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                                   __uint_as_float(r[4 * j + 3]));
          }
          if (chunk == 1 && head_ok) {
            uint32_t ks;
            tmem_ld1(kv_addr + 32, ks);
            tmem_ld_wait();
            prow[32 * 32 + lane] = __uint_as_float(ks);              // Ks_head[dk = lane]
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_remote(kvempty_leader);
        }
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();                                // nobody exits (or frees TMEM) while the peer still needs this CTA
  if (warp == 2) tmem_dealloc2(tmem, 512);
}

}  // namespace
}  // namespace yat

extern "C" int64_t yat_qkv_linattn_state_bytes(int B, int N, int H) {
  const int64_t m_blocks = (N + 255) / 256;
  return (int64_t)B * 2 * m_blocks * H * yat::kStateRow * (int64_t)sizeof(float);
}

extern "C" int yat_qkv_linattn_fwd(const void* x, const void* w_qkv, int B, int N, int H, int D, int Cin, void* qkv,
                                   void* state, void* stream) {
  using namespace yat;
  YAT_REQUIRE(x && w_qkv && qkv && state, YAT_ERR_BAD_ARG, "yat_qkv_linattn_fwd: null pointer argument");
  YAT_REQUIRE(B > 0 && N > 0 && H > 0 && Cin > 0, YAT_ERR_BAD_ARG, "yat_qkv_linattn_fwd: non-positive dimension");
  YAT_REQUIRE(D == 32, YAT_ERR_UNSUPPORTED, "yat_qkv_linattn_fwd: head_dim %d not supported (SANA uses 32)", D);
  YAT_REQUIRE(Cin % 8 == 0, YAT_ERR_BAD_ARG, "yat_qkv_linattn_fwd: Cin must be a multiple of 8 (16-byte TMA rows)");
  YAT_REQUIRE((((uintptr_t)x | (uintptr_t)w_qkv | (uintptr_t)state) & 15u) == 0 && ((uintptr_t)qkv & 31u) == 0, YAT_ERR_BAD_ARG,
              "yat_qkv_linattn_fwd: x / w / state must be 16-byte aligned, qkv 32-byte");
  const int C = H * D;
  QkvMaps maps;
  int rc;
  if ((rc = make_tmap_3d_bf16(&maps.x, x, (uint64_t)Cin, (uint64_t)N, (uint64_t)B, 64, kBM)) != YAT_OK) return rc;
  if ((rc = make_tmap_2d_bf16(&maps.w, w_qkv, (uint64_t)Cin, (uint64_t)3 * C, (uint64_t)Cin, kHalfN)) != YAT_OK) return rc;
  QkvArgs args{(__nv_bfloat16*)qkv, (float*)state, B, N, H, C, Cin};
  static bool attr_set = false;
  if (!attr_set) {
    const cudaError_t e = cudaFuncSetAttribute(qkv_linattn_sm100_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kQkvSmemBytes);
    YAT_REQUIRE(e == cudaSuccess, YAT_ERR_CUDA, "yat_qkv_linattn_fwd: cannot reserve %d B of shared memory (%s)", kQkvSmemBytes,
                cudaGetErrorString(e));
    attr_set = true;
  }
  const int m_blocks = (N + 255) / 256;
  const int tiles = B * m_blocks * ((C + kBN - 1) / kBN + (H + kGroupHeads - 1) / kGroupHeads);
  const int clusters = tiles < kNumSMs / 2 ? tiles : kNumSMs / 2;
  qkv_linattn_sm100_kernel<<<2 * clusters, kThreads, kQkvSmemBytes, (cudaStream_t)stream>>>(maps, args);
  return check_launch("qkv_linattn_sm100");
}
