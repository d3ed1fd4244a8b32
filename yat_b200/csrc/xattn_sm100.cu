// Cross-attention of the SANA / PixArt blocks (attn2) on the 5th-gen tensor cores (tcgen05 + TMEM + TMA), forward.
//
//   O_i = sum_j softmax_j(scale q_i.k_j + bias_j) v_j            per (batch, head);  bias_j = 0 | -10000 (text padding)
//
// replaces F.scaled_dot_product_attention inside diffusers' AttnProcessor2_0 for the block's attn2 (reached from
// utils/patch_sana_attention_layers.py:100-106), which the library runs at head dim 128 on padded copies of q and k.
//
// Shape of the problem: many queries (1024+ image tokens), FEW keys (a caption: L <= 384 after trimming), head dim
// d <= 128 (SANA-1.6B: 112).  A whole row of scores therefore fits in tensor memory: S = Q K^T is 128 lanes x L fp32
// columns next to the 128 columns of O, and the softmax is exact two-pass (row max, then exp / sum) with no online
// rescaling.  Q, K, V and O stay token-major [B, rows, H*d]: a 4-D tensor map {d, H, rows, B} addresses one head, and the
// columns d..127 of a 64-wide box come back as zeros (and are clipped on store), so d = 112 costs no padded copies.
//
// Work item = (batch, head, 128-query tile); persistent grid, one CTA per SM.  Per item:
//   producer   Q tile (2 x [128 x 64] SW128 boxes), then the K chunks and the V chunks ([64 keys x 128] each) through a ring
//   issuer     S[:, 64c..] = Q K_c^T (K-major A and B)  ->  s_full;   O += P_c V_c (A = P_c written by the softmax warps,
//              B = V_c MN-major straight from its TMA tile), chunk by chunk as P arrives  ->  o_full
//   softmax    8 warps = 2 per TMEM lane quadrant (thread = query row x half of each chunk's columns): pass 1 row max of
//              s*c + bias2, pass 2 p = 2^(s*c + bias2 - m) -> bf16 -> swizzled K-major P chunk in shared memory, row sum;
//              epilogue O / l -> bf16 -> the item's Q buffer -> TMA store; lse = ln sum_j exp(...) for the backward
// c = scale * log2(e), bias2 = bias * log2(e) (keys past L: -inf).
#include "common.cuh"
#include "sm100.cuh"

namespace yat {
namespace {
using namespace sm100;

constexpr int kQT = 128;                     // queries per tile
constexpr int kKC = 64;                      // keys per chunk
constexpr int kMaxChunks = 6;                // L <= 384: S occupies TMEM columns [0, 384), O [384, 512)
constexpr int kQHalf = kQT * 128;            // [128 rows][64 cols] bf16
constexpr int kQBytes = 2 * kQHalf;          // 32 KB
constexpr int kCHalf = kKC * 128;            // [64 keys][64 cols] bf16
constexpr int kCBytes = 2 * kCHalf;          // 16 KB
constexpr int kPBytes = kQT * 128;           // [128 rows][64 keys] bf16
constexpr int kStages = 6;
constexpr int kSoftWarps = 8, kSoftThreads = 32 * kSoftWarps;
constexpr int kStoreWarp = 2 + kSoftWarps;
constexpr int kThreads = 32 * (kStoreWarp + 1);
constexpr uint32_t kColO = 64 * kMaxChunks;
constexpr int kTmemCols = 512;

struct XfSmem {
  uint8_t q[2][kQBytes];                     // 1024-aligned; the O tile leaves from the item's Q buffer
  uint8_t kv[kStages][kCBytes];
  uint8_t p[2][kPBytes];
  float bias2[2][kMaxChunks * kKC];
  float xm[2][kQT], xl[2][kQT];              // row max / row sum halves exchanged between the two warps of a row
  uint64_t q_full[2], q_empty[2], kv_full[kStages], kv_empty[kStages], p_ready[2], p_empty[2];
  uint64_t s_full, s_empty, o_full, o_empty, o_staged;
  uint32_t tmem_base;
};
constexpr int kXfSmemBytes = (int)sizeof(XfSmem) + 1024;
static_assert(kXfSmemBytes <= 232448, "cross-attention forward shared memory exceeds the 227 KB per-CTA limit");

constexpr uint32_t kIdescS = make_idesc_bf16(128, kKC, 0, 0);    // Q K_c^T : A, B K-major
constexpr uint32_t kIdescPV = make_idesc_bf16(128, 128, 0, 1);   // P_c V_c : A K-major, B MN-major

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct Item { int b, h, q0; };
__device__ __forceinline__ Item decode(int item, int n_qt, int H) {
  Item r;
  r.q0 = (item % n_qt) * kQT;
  const int bh = item / n_qt;
  r.h = bh % H;
  r.b = bh / H;
  return r;
}

__global__ void __launch_bounds__(kThreads, 1)
xattn_fwd_sm100_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_o,
                       const float* __restrict__ bias, float* __restrict__ lse, int B, int H, int Nq, int L, int d,
                       float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  XfSmem& S = *reinterpret_cast<XfSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_qt = (Nq + kQT - 1) / kQT;
  const int n_items = B * H * n_qt;
  const int nC = (L + kKC - 1) / kKC;          // key chunks
  const int nK = (d + 15) / 16;                // K steps of the score MMA (columns d.. of Q and K are zero-filled)

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&S.q_full[i]), 1);
      mbar_init(smem_u32(&S.q_empty[i]), 1);
      mbar_init(smem_u32(&S.p_ready[i]), kSoftWarps);
      mbar_init(smem_u32(&S.p_empty[i]), 1);
    }
    for (int s = 0; s < kStages; ++s) {
      mbar_init(smem_u32(&S.kv_full[s]), 1);
      mbar_init(smem_u32(&S.kv_empty[s]), 1);
    }
    mbar_init(smem_u32(&S.s_full), 1);
    mbar_init(smem_u32(&S.s_empty), kSoftWarps);
    mbar_init(smem_u32(&S.o_full), 1);
    mbar_init(smem_u32(&S.o_empty), kSoftWarps);
    mbar_init(smem_u32(&S.o_staged), kSoftWarps);
    fence_barrier_init();
    tma_prefetch_desc(&tm_q); tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v); tma_prefetch_desc(&tm_o);
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(&S.tmem_base), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (S.tmem_base != 0) __trap();              // one CTA per SM allocating all 512 columns: the base is column 0
  constexpr uint32_t tmem = 0;

  if (warp == 0) {
    // ======================= TMA producer =======================
    if (lane == 0) {
      uint32_t it = 0, qi = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi) {
        const Item w = decode(item, n_qt, H);
        const uint32_t qb = qi & 1;
        mbar_wait(smem_u32(&S.q_empty[qb]), ((qi >> 1) & 1) ^ 1);
        const uint32_t qf = smem_u32(&S.q_full[qb]);
        mbar_arrive_expect_tx(qf, kQBytes);
        tma_load_4d(smem_u32(&S.q[qb][0]), &tm_q, qf, 0, w.h, w.q0, w.b);
        tma_load_4d(smem_u32(&S.q[qb][kQHalf]), &tm_q, qf, 64, w.h, w.q0, w.b);
        for (int j = 0; j < 2 * nC; ++j, ++it) {
          const CUtensorMap* tm = j < nC ? &tm_k : &tm_v;
          const int c = j < nC ? j : j - nC;
          const uint32_t s = it % kStages, ph = (it / kStages) & 1;
          mbar_wait(smem_u32(&S.kv_empty[s]), ph ^ 1);
          const uint32_t full = smem_u32(&S.kv_full[s]);
          mbar_arrive_expect_tx(full, kCBytes);
          tma_load_4d(smem_u32(&S.kv[s][0]), tm, full, 0, w.h, c * kKC, w.b);
          tma_load_4d(smem_u32(&S.kv[s][kCHalf]), tm, full, 64, w.h, c * kKC, w.b);
        }
      }
    }
  } else if (warp == 1) {
    // ======================= tcgen05 issuer =======================
    if (lane == 0) {
      uint32_t it = 0, qi = 0, p_use = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi) {
        const uint32_t qb = qi & 1;
        mbar_wait(smem_u32(&S.q_full[qb]), (qi >> 1) & 1);
        mbar_wait(smem_u32(&S.s_empty), (qi & 1) ^ 1);       // the previous item's scores have been read
        tc_fence_after();
        // K-major SW128 operands: 8-row groups 1 KB apart, K step of 16 = 32 B inside the 128 B row, 64-column halves
        const uint64_t qa0 = make_smem_desc(smem_u32(&S.q[qb][0]), 16, 1024, kLayoutSw128);
        for (int c = 0; c < nC; ++c, ++it) {
          const uint32_t s = it % kStages;
          mbar_wait(smem_u32(&S.kv_full[s]), (it / kStages) & 1);
          tc_fence_after();
          const uint64_t kb0 = make_smem_desc(smem_u32(&S.kv[s][0]), 16, 1024, kLayoutSw128);
          for (int ks = 0; ks < nK; ++ks)
            umma_bf16(tmem + c * kKC, qa0 + (((ks >> 2) * kQHalf + (ks & 3) * 32) >> 4),
                      kb0 + (((ks >> 2) * kCHalf + (ks & 3) * 32) >> 4), kIdescS, ks != 0);
          tc_commit(smem_u32(&S.kv_empty[s]));
        }
        tc_commit(smem_u32(&S.s_full));
        for (int c = 0; c < nC; ++c, ++it, ++p_use) {
          const uint32_t s = it % kStages, pb = p_use & 1;
          mbar_wait(smem_u32(&S.kv_full[s]), (it / kStages) & 1);
          mbar_wait(smem_u32(&S.p_ready[pb]), (p_use >> 1) & 1);
          if (c == 0) mbar_wait(smem_u32(&S.o_empty), (qi & 1) ^ 1);   // the previous item's O has left TMEM
          tc_fence_after();
          const uint64_t pa0 = make_smem_desc(smem_u32(&S.p[pb][0]), 16, 1024, kLayoutSw128);
          // V_c as the B operand, MN-major SW128: LBO = stride between the two 64-column halves, SBO = 8-key group (1 KB),
          // K step of 16 keys = 2 KB
          const uint64_t vb0 = make_smem_desc(smem_u32(&S.kv[s][0]), kCHalf, 1024, kLayoutSw128);
#pragma unroll
          for (int ks = 0; ks < kKC / 16; ++ks)
            umma_bf16(tmem + kColO, pa0 + ((ks * 32) >> 4), vb0 + ((ks * 2048) >> 4), kIdescPV, (c | ks) != 0);
          tc_commit(smem_u32(&S.kv_empty[s]));
          tc_commit(smem_u32(&S.p_empty[pb]));
        }
        tc_commit(smem_u32(&S.o_full));
      }
    }
  } else if (warp == kStoreWarp) {
    // ======================= TMA-store warp =======================
    if (lane == 0) {
      uint32_t qi = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi) {
        const Item w = decode(item, n_qt, H);
        const uint32_t qb = qi & 1;
        mbar_wait(smem_u32(&S.o_staged), qi & 1);
        tma_store_4d(&tm_o, smem_u32(&S.q[qb][0]), 0, w.h, w.q0, w.b);
        if (d > 64) tma_store_4d(&tm_o, smem_u32(&S.q[qb][kQHalf]), 64, w.h, w.q0, w.b);
        tma_store_commit();
        tma_store_wait_read<0>();                 // shared memory has been read: the Q buffer can be refilled
        mbar_arrive(smem_u32(&S.q_empty[qb]));
      }
      tma_store_wait_all<0>();
    }
  } else {
    // ======================= softmax + epilogue warps =======================
    const int te = threadIdx.x - 64;              // 0..255
    const int wg = te >> 7;                       // which 32 of a chunk's 64 columns / which half of O
    const int quad = warp & 3;                    // TMEM lane quadrant this warp may access
    const int row = quad * 32 + lane;             // query row in the tile (TMEM lane)
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const f2 cs2 = bc2(scale_log2);
    uint32_t qi = 0, p_use = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi) {
      const Item w = decode(item, n_qt, H);
      const uint32_t qb = qi & 1;
      float* bs = S.bias2[qi & 1];
      for (int j = te; j < nC * kKC; j += kSoftThreads)
        bs[j] = j < L ? (bias ? __ldg(bias + (int64_t)w.b * L + j) * 1.4426950408889634f : 0.f) : -INFINITY;
      named_bar_sync(1, kSoftThreads);
      mbar_wait(smem_u32(&S.s_full), qi & 1);
      tc_fence_after();
      // ---- pass 1: row max of s * c + bias2 over this thread's half of every chunk
      float mx = -INFINITY;
      for (int c = 0; c < nC; ++c) {
        uint32_t r[32];
        tmem_ld32(tmem + lane_base + c * kKC + wg * 32, r);
        tmem_ld_wait();
        const float4* b4 = reinterpret_cast<const float4*>(bs + c * kKC + wg * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 bb = b4[j];
          float a0, a1, a2, a3;
          un2(fma2(mk2(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1])), cs2, mk2(bb.x, bb.y)), a0, a1);
          un2(fma2(mk2(__uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3])), cs2, mk2(bb.z, bb.w)), a2, a3);
          mx = fmaxf(fmaxf(mx, fmaxf(a0, a1)), fmaxf(a2, a3));
        }
      }
      S.xm[wg][row] = mx;
      named_bar_sync(1, kSoftThreads);
      mx = fmaxf(mx, S.xm[wg ^ 1][row]);
      // ---- pass 2: p = 2^(s c + bias2 - m) -> bf16 P chunk (K-major SW128: row = 128 B, 16 B chunks XOR-swizzled)
      const f2 nm = bc2(-mx);
      float sum = 0.f;
      for (int c = 0; c < nC; ++c, ++p_use) {
        const uint32_t pb = p_use & 1;
        uint32_t r[32];
        tmem_ld32(tmem + lane_base + c * kKC + wg * 32, r);
        mbar_wait(smem_u32(&S.p_empty[pb]), ((p_use >> 1) & 1) ^ 1);
        tmem_ld_wait();
        const float4* b4 = reinterpret_cast<const float4*>(bs + c * kKC + wg * 32);
        uint8_t* prow = &S.p[pb][row * 128];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t wv[4];
#pragma unroll
          for (int jj = 0; jj < 2; ++jj) {
            const int j = 2 * g + jj;
            const float4 bb = b4[j];
            float a0, a1, a2, a3;
            un2(add2(fma2(mk2(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1])), cs2, mk2(bb.x, bb.y)), nm), a0, a1);
            un2(add2(fma2(mk2(__uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3])), cs2, mk2(bb.z, bb.w)), nm), a2, a3);
            const float p0 = ex2(a0), p1 = ex2(a1), p2 = ex2(a2), p3 = ex2(a3);
            sum += (p0 + p1) + (p2 + p3);
            wv[2 * jj] = pack_bf16(p0, p1);
            wv[2 * jj + 1] = pack_bf16(p2, p3);
          }
          *reinterpret_cast<uint4*>(prow + (((wg * 4 + g) ^ (row & 7)) << 4)) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.p_ready[pb]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&S.s_empty));
      S.xl[wg][row] = sum;
      named_bar_sync(1, kSoftThreads);
      sum += S.xl[wg ^ 1][row];
      // ---- epilogue: O / l -> bf16 -> this item's Q buffer (half wg) -> store warp
      mbar_wait(smem_u32(&S.o_full), qi & 1);
      tc_fence_after();
      const float rinv = __frcp_rn(sum);
      uint8_t* orow = &S.q[qb][wg * kQHalf + row * 128];
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        uint32_t r[32];
        tmem_ld32(tmem + lane_base + kColO + wg * 64 + hh * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t wv[4];
#pragma unroll
          for (int j = 0; j < 4; ++j)
            wv[j] = pack_bf16(__uint_as_float(r[g * 8 + 2 * j]) * rinv, __uint_as_float(r[g * 8 + 2 * j + 1]) * rinv);
          *reinterpret_cast<uint4*>(orow + (((hh * 4 + g) ^ (row & 7)) << 4)) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
        }
      }
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(smem_u32(&S.o_empty));
        mbar_arrive(smem_u32(&S.o_staged));
      }
      if (wg == 0 && lse != nullptr && w.q0 + row < Nq)
        lse[((int64_t)w.b * H + w.h) * Nq + w.q0 + row] = 0.6931471805599453f * (mx + log2f(sum));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

}  // namespace
}  // namespace yat

// ------------------------------------------------------------------ host side
extern "C" int yat_xattn_max_keys() { return yat::kMaxChunks * yat::kKC; }

extern "C" int yat_xattn_fwd(const void* q, const void* k, const void* v, const void* bias, int B, int H, int Nq, int L, int d,
                             int64_t ldq, int64_t ldkv, float scale, void* o, void* lse, void* stream) {
  using namespace yat;
  YAT_REQUIRE(q && k && v && o, YAT_ERR_BAD_ARG, "yat_xattn_fwd: null pointer argument");
  YAT_REQUIRE(B > 0 && H > 0 && Nq > 0 && L > 0, YAT_ERR_BAD_ARG, "yat_xattn_fwd: non-positive dimension");
  YAT_REQUIRE(d >= 8 && d <= 128 && d % 8 == 0, YAT_ERR_UNSUPPORTED, "yat_xattn_fwd: head_dim %d not supported (8..128, multiple of 8)", d);
  YAT_REQUIRE(L <= kMaxChunks * kKC, YAT_ERR_UNSUPPORTED, "yat_xattn_fwd: %d keys exceed the %d that fit in tensor memory", L,
              kMaxChunks * kKC);
  YAT_REQUIRE(ldq >= (int64_t)H * d && ldq % 8 == 0 && ldkv >= (int64_t)H * d && ldkv % 8 == 0, YAT_ERR_BAD_ARG,
              "yat_xattn_fwd: row pitches %lld / %lld must be >= H*d and multiples of 8", (long long)ldq, (long long)ldkv);
  YAT_REQUIRE((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)o) & 15u) == 0, YAT_ERR_BAD_ARG,
              "yat_xattn_fwd: tensors must be 16-byte aligned");
  CUtensorMap tq, tk, tv, to;
  int rc;
  if ((rc = make_tmap_heads_bf16(&tq, q, d, H, Nq, B, (uint64_t)ldq, kQT)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tk, k, d, H, L, B, (uint64_t)ldkv, kKC)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tv, v, d, H, L, B, (uint64_t)ldkv, kKC)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&to, o, d, H, Nq, B, (uint64_t)H * d, kQT)) != YAT_OK) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(xattn_fwd_sm100_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kXfSmemBytes);
    YAT_REQUIRE(e == cudaSuccess, YAT_ERR_CUDA, "yat_xattn_fwd: cannot reserve %d B of shared memory (%s)", kXfSmemBytes,
                cudaGetErrorString(e));
    attr_set = true;
  }
  const int items = B * H * ((Nq + kQT - 1) / kQT);
  const int grid = items < kNumSMs ? items : kNumSMs;
  xattn_fwd_sm100_kernel<<<grid, kThreads, kXfSmemBytes, (cudaStream_t)stream>>>(
      tq, tk, tv, to, (const float*)bias, (float*)lse, B, H, Nq, L, d, scale * 1.4426950408889634f);
  return check_launch("xattn_fwd_sm100");
}
