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
// Work item = (batch, head, 128-query tile); persistent grid, one CTA per SM.  The probabilities never touch shared
// memory: the softmax threads write P (bf16 pairs) back into the tensor-memory columns of the scores they came from, and
// P_c V_c reads its A operand from tensor memory.
//   producer   Q tile (2 x [128 x 64] SW128 boxes); K and V chunks ([128 keys x 128]) through one ring, in the order the issuer
//              consumes them: K_0.. of the CTA's first item, then per item i: V_c(i) with the early K_c(i+1) behind them, the
//              last two K chunks of item i+1 at the end
//   issuer     O += P_j V_j per 64-key sub-chunk (A = P_j in tensor memory, B = V MN-major straight from its TMA tile) and,
//              behind the second sub-chunk of a ring chunk, S'[:, 128c..] = Q' K'_c^T of the NEXT item (N = 128: a 64-key
//              score MMA is bound by its shared-memory operand reads) into the score columns pass 2 has just finished with:
//              the next item's scores are complete when this item's pass 2 ends
//   softmax    8 warps = 2 per TMEM lane quadrant (thread = query row x half of each chunk's columns): pass 1 row max of
//              s*c + bias2, pass 2 p = 2^(s*c + bias2 - m) -> bf16 pairs -> tcgen05.st over the thread's own score columns,
//              row sum; epilogue O / l -> bf16 -> the item's Q buffer -> TMA store; lse = ln sum_j exp(...) for a backward
// c = scale * log2(e), bias2 = bias * log2(e) (keys past L: -inf).
#include "common.cuh"
#include "sm100.cuh"

namespace yat {
namespace {
using namespace sm100;

constexpr int kQT = 128;                     // queries per tile
constexpr int kKC = 64;                      // keys per softmax sub-chunk (one p_ready barrier, 4 P V MMAs)
constexpr int kKR = 128;                     // keys per ring chunk (one TMA pair, one N = 128 score MMA per K step)
constexpr int kMaxChunks = 6;                // L <= 384: S occupies TMEM columns [0, 384), O [384, 512)
constexpr int kQHalf = kQT * 128;            // [128 rows][64 cols] bf16
constexpr int kQBytes = 2 * kQHalf;          // 32 KB
constexpr int kCHalf = kKR * 128;            // [128 keys][64 cols] bf16
constexpr int kCBytes = 2 * kCHalf;          // 32 KB
constexpr int kStages = 4;
constexpr int kQBufs = 2;
constexpr int kSoftWarps = 8, kSoftThreads = 32 * kSoftWarps;
constexpr int kStoreWarp = 2 + kSoftWarps;
constexpr int kThreads = 32 * (kStoreWarp + 1);
constexpr uint32_t kColO = 64 * kMaxChunks;
constexpr int kTmemCols = 512;

struct XfSmem {
  uint8_t q[kQBufs][kQBytes];                // 1024-aligned; the O tile leaves from the item's Q buffer
  uint8_t kv[kStages][kCBytes];
  float bias2[2][kMaxChunks * kKC];
  float xm[2][kQT], xl[2][kQT];              // row max / row sum halves exchanged between the two warps of a row
  uint64_t q_full[kQBufs], kv_full[kStages], kv_empty[kStages], p_ready[kMaxChunks];
  uint64_t s_full, o_full, o_empty, o_staged;
  uint32_t tmem_base;
};
constexpr int kXfSmemBytes = (int)sizeof(XfSmem) + 1024;
static_assert(kXfSmemBytes <= 232448, "cross-attention forward shared memory exceeds the 227 KB per-CTA limit");

constexpr uint32_t kIdescS128 = make_idesc_bf16(128, 128, 0, 0);  // Q K_c^T over 128 keys: A, B K-major
constexpr uint32_t kIdescS64 = make_idesc_bf16(128, 64, 0, 0);    // last chunk with <= 64 keys
// P_j V_j : A K-major (tensor memory), B MN-major; N = head dim rounded up to 16 (112 for SANA-1.6B: an eighth less work)
__device__ __forceinline__ uint32_t idesc_pv(int n) { return make_idesc_bf16(128, n, 0, 1); }

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
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
  const int nC = (L + kKC - 1) / kKC;          // 64-key sub-chunks
  const int nR = (L + kKR - 1) / kKR;          // 128-key ring chunks
  const int nK = (d + 15) / 16;                // K steps of the score MMA (columns d.. of Q and K are zero-filled)

  if (threadIdx.x == 0) {
    for (int i = 0; i < kQBufs; ++i) {
      mbar_init(smem_u32(&S.q_full[i]), 1);
    }
    for (int s = 0; s < kStages; ++s) {
      mbar_init(smem_u32(&S.kv_full[s]), 1);
      mbar_init(smem_u32(&S.kv_empty[s]), 1);
    }
    for (int c = 0; c < kMaxChunks; ++c) mbar_init(smem_u32(&S.p_ready[c]), kSoftWarps);
    mbar_init(smem_u32(&S.s_full), 1);
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
      auto load_chunk = [&](const CUtensorMap* tm, const Item& w, int c) {
        const uint32_t s = it % kStages, ph = (it / kStages) & 1;
        mbar_wait(smem_u32(&S.kv_empty[s]), ph ^ 1);
        const uint32_t full = smem_u32(&S.kv_full[s]);
        mbar_arrive_expect_tx(full, kCBytes);
        tma_load_4d(smem_u32(&S.kv[s][0]), tm, full, 0, w.h, c * kKR, w.b);
        tma_load_4d(smem_u32(&S.kv[s][kCHalf]), tm, full, 64, w.h, c * kKR, w.b);
        ++it;
      };
      if ((int)blockIdx.x < n_items) {
        const Item w0 = decode(blockIdx.x, n_qt, H);
        for (int c = 0; c < nR; ++c) load_chunk(&tm_k, w0, c);
      }
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi) {
        const Item w = decode(item, n_qt, H);
        const bool has_next = item + (int)gridDim.x < n_items;
        Item wn = w;
        if (has_next) wn = decode(item + gridDim.x, n_qt, H);
        for (int c = 0; c < nR; ++c) {               // the order the issuer consumes them in (see there)
          load_chunk(&tm_v, w, c);
          if (has_next && c < nR - 2) load_chunk(&tm_k, wn, c);
        }
        if (has_next)
          for (int c = max(0, nR - 2); c < nR; ++c) load_chunk(&tm_k, wn, c);
      }
    }
  } else if (warp == 1) {
    // ======================= tcgen05 issuer =======================
    if (lane == 0 && (int)blockIdx.x < n_items) {
      uint32_t it = 0, qi = 0;
      const uint32_t idesc_o = idesc_pv((d + 15) & ~15);
      // scores of ring chunk c (128 keys, or 64 when no more are left) of the item in Q buffer qb: K-major SW128 operands
      // (8-row groups 1 KB apart, K step of 16 = 32 B inside the 128 B row, 64-column halves)
      auto score_mma = [&](uint32_t qb, int c) {
        const uint32_t s = it % kStages;
        mbar_wait(smem_u32(&S.kv_full[s]), (it / kStages) & 1);
        tc_fence_after();
        const uint32_t idesc = (nC - 2 * c >= 2) ? kIdescS128 : kIdescS64;
        const uint64_t qa0 = make_smem_desc(smem_u32(&S.q[qb][0]), 16, 1024, kLayoutSw128);
        const uint64_t kb0 = make_smem_desc(smem_u32(&S.kv[s][0]), 16, 1024, kLayoutSw128);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          if (ks < nK)
            umma_bf16(tmem + c * kKR, qa0 + (((ks >> 2) * kQHalf + (ks & 3) * 32) >> 4),
                      kb0 + (((ks >> 2) * kCHalf + (ks & 3) * 32) >> 4), idesc, ks != 0);
        tc_commit(smem_u32(&S.kv_empty[s]));
        ++it;
      };
      mbar_wait(smem_u32(&S.q_full[0]), 0);
      for (int c = 0; c < nR; ++c) score_mma(0, c);
      tc_commit(smem_u32(&S.s_full));
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi) {
        const bool has_next = item + (int)gridDim.x < n_items;
#ifdef YAT_BWD_PROBE
        long long t_v = 0, t_p = 0, t_pv = 0, t_s = 0;   // cycles: waiting for V, for P, issuing P V, issuing the next scores
#define XT(acc, ...) do { const long long t0_ = clock64(); __VA_ARGS__; acc += clock64() - t0_; } while (0)
#else
#define XT(acc, ...) do { __VA_ARGS__; } while (0)
#endif
        for (int j = 0; j < nC; ++j) {
          const int c = j >> 1;
          const uint32_t s = it % kStages;
          if ((j & 1) == 0) XT(t_v, mbar_wait(smem_u32(&S.kv_full[s]), (it / kStages) & 1));
          XT(t_p, mbar_wait(smem_u32(&S.p_ready[j]), qi & 1);                  // every softmax warp has written P_j over S_j
             if (j == 0) mbar_wait(smem_u32(&S.o_empty), (qi & 1) ^ 1));      // the previous item's O has left TMEM
          tc_fence_after();
          // V as the B operand, MN-major SW128: LBO = stride between the two 64-column halves, SBO = 8-key group (1 KB),
          // K step of 16 keys = 2 KB.  A = P_j in tensor memory: keys 0..31 of the sub-chunk as bf16 pairs in columns
          // [64j, 64j+16), keys 32..63 in [64j+32, 64j+48) (each softmax thread wrote over its own score columns)
          const uint64_t vb0 = make_smem_desc(smem_u32(&S.kv[s][0]), kCHalf, 1024, kLayoutSw128);
          const bool last_of_chunk = (j & 1) == 1 || j == nC - 1;
          auto pv_mma = [&]() {
#pragma unroll
            for (int ks = 0; ks < kKC / 16; ++ks)
              umma_bf16_ts(tmem + kColO, tmem + j * kKC + (ks >> 1) * 32 + (ks & 1) * 8,
                           vb0 + ((((j & 1) * 4 + ks) * 2048) >> 4), idesc_o, (j | ks) != 0);
            if (last_of_chunk) { tc_commit(smem_u32(&S.kv_empty[s])); ++it; }
            if (j == nC - 1) tc_commit(smem_u32(&S.o_full));
          };
          XT(t_pv, pv_mma());
          // the next item's scores go into the columns of ring chunks whose P V has been issued (the tensor pipe runs in
          // order).  Early chunks are issued right away; the last two wait until the final P V is in the pipe, so that
          // nothing sits between the end of pass 2 and o_full - they run under the epilogue instead.
          if (has_next && last_of_chunk) {
            const uint32_t qn = (qi + 1) & 1;
            const int c_lo = (j == nC - 1) ? max(0, min(c, nR - 2)) : c, c_hi = (j == nC - 1 || c < nR - 2) ? c : c_lo - 1;
            XT(t_s,
            for (int cc = c_lo; cc <= c_hi; ++cc) {
              if (cc == 0) mbar_wait(smem_u32(&S.q_full[qn]), ((qi + 1) >> 1) & 1);
              score_mma(qn, cc);
            });
          }
        }
#ifdef YAT_BWD_PROBE
        TRACE(1, 1000 + (int)(t_v >> 4)); TRACE(1, 2000 + (int)(t_p >> 4)); TRACE(1, 3000 + (int)(t_pv >> 4)); TRACE(1, 4000 + (int)(t_s >> 4));
#endif
        if (has_next) tc_commit(smem_u32(&S.s_full));
      }
    }
  } else if (warp == kStoreWarp) {
    // ======================= Q loads + O stores (one thread: a Q buffer is refilled right after its O tile has left) ==========
    if (lane == 0) {
      auto load_q = [&](int item, uint32_t qb) {
        const Item w = decode(item, n_qt, H);
        const uint32_t qf = smem_u32(&S.q_full[qb]);
        mbar_arrive_expect_tx(qf, kQBytes);
        tma_load_4d(smem_u32(&S.q[qb][0]), &tm_q, qf, 0, w.h, w.q0, w.b);
        tma_load_4d(smem_u32(&S.q[qb][kQHalf]), &tm_q, qf, 64, w.h, w.q0, w.b);
      };
      if ((int)blockIdx.x < n_items) load_q(blockIdx.x, 0);
      if ((int)(blockIdx.x + gridDim.x) < n_items) load_q(blockIdx.x + gridDim.x, 1);
      uint32_t qi = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi) {
        const Item w = decode(item, n_qt, H);
        const uint32_t qb = qi & 1;
        mbar_wait(smem_u32(&S.o_staged), qi & 1);
        tma_store_4d(&tm_o, smem_u32(&S.q[qb][0]), 0, w.h, w.q0, w.b);
        if (d > 64) tma_store_4d(&tm_o, smem_u32(&S.q[qb][kQHalf]), 64, w.h, w.q0, w.b);
        tma_store_commit();
        tma_store_wait_read<0>();                 // shared memory has been read: the buffer takes the Q tile two items ahead
        if (item + 2 * (int)gridDim.x < n_items) load_q(item + 2 * gridDim.x, qb);
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
    const uint32_t s_addr = tmem + lane_base + wg * 32;   // this thread's 32 score columns of chunk 0 (P goes to the first 16)
    const f2 cs2 = bc2(scale_log2);
    constexpr float kLog2e = 1.4426950408889634f;
    uint32_t qi = 0;
    if ((int)blockIdx.x < n_items) {
      const Item w0 = decode(blockIdx.x, n_qt, H);
      for (int j = te; j < nC * kKC; j += kSoftThreads)
        S.bias2[0][j] = j < L ? (bias ? __ldg(bias + (int64_t)w0.b * L + j) * kLog2e : 0.f) : -INFINITY;
      named_bar_sync(1, kSoftThreads);
    }
    Item w_next = decode(blockIdx.x, n_qt, H);
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi) {
      const Item w = w_next;
      const bool has_next = item + (int)gridDim.x < n_items;
      const float* bs = S.bias2[qi & 1];
      float nb0 = 0.f, nb1 = 0.f;                 // the next item's bias row: in flight during pass 1
      if (has_next) {
        w_next = decode(item + gridDim.x, n_qt, H);
        if (bias != nullptr) {
          const float* bn = bias + (int64_t)w_next.b * L;
          if (te < L) nb0 = __ldg(bn + te);
          if (te + kSoftThreads < L) nb1 = __ldg(bn + te + kSoftThreads);
        }
      }
      if (te == 0) TRACE(0, 10);
      mbar_wait(smem_u32(&S.s_full), qi & 1);
      tc_fence_after();
      if (te == 0) TRACE(0, 11);
      // ---- pass 1: row max of s * c + bias2 over this thread's half of every chunk (the load of chunk c+1 is in flight)
      float mx = -INFINITY;
      auto max_chunk = [&](const uint32_t (&r)[32], int c) {
        const float4* b4 = reinterpret_cast<const float4*>(bs + c * kKC + wg * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 bb = b4[j];
          float a0, a1, a2, a3;
          un2(fma2(mk2(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1])), cs2, mk2(bb.x, bb.y)), a0, a1);
          un2(fma2(mk2(__uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3])), cs2, mk2(bb.z, bb.w)), a2, a3);
          mx = max3(max3(mx, a0, a1), a2, a3);
        }
      };
      {
        uint32_t ra[32], rb[32];
        tmem_ld32(s_addr, ra);
        for (int c = 0; c < nC; c += 2) {
          tmem_ld_wait();
          if (c + 1 < nC) tmem_ld32(s_addr + (c + 1) * kKC, rb);
          max_chunk(ra, c);
          if (c + 1 < nC) {
            tmem_ld_wait();
            if (c + 2 < nC) tmem_ld32(s_addr + (c + 2) * kKC, ra);
            max_chunk(rb, c + 1);
          }
        }
      }
      if (te == 0) TRACE(0, 12);
      S.xm[wg][row] = mx;
      if (has_next) {      // next item's bias2 (that buffer was last read in the previous item's pass 2: two barriers ago)
        float* bn = S.bias2[(qi + 1) & 1];
        if (te < nC * kKC) bn[te] = te < L ? nb0 * kLog2e : -INFINITY;
        if (te + kSoftThreads < nC * kKC) bn[te + kSoftThreads] = te + kSoftThreads < L ? nb1 * kLog2e : -INFINITY;
      }
      named_bar_sync(1, kSoftThreads);
      if (te == 0) TRACE(0, 13);
      mx = fmaxf(mx, S.xm[wg ^ 1][row]);
      // ---- pass 2: p = 2^(s c + bias2 - m) -> bf16 pairs -> back into the first 16 of this thread's 32 score columns
      const f2 nm = bc2(-mx);
      float sum = 0.f;
      auto exp_chunk = [&](const uint32_t (&r)[32], int c) {
        const float4* b4 = reinterpret_cast<const float4*>(bs + c * kKC + wg * 32);
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 bb = b4[j];
          float a0, a1, a2, a3;
          un2(add2(fma2(mk2(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1])), cs2, mk2(bb.x, bb.y)), nm), a0, a1);
          un2(add2(fma2(mk2(__uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3])), cs2, mk2(bb.z, bb.w)), nm), a2, a3);
          const float p0 = ex2(a0), p1 = ex2(a1), p2 = ex2(a2), p3 = ex2(a3);
          sum += (p0 + p1) + (p2 + p3);
          pk[2 * j] = pack_bf16(p0, p1);
          pk[2 * j + 1] = pack_bf16(p2, p3);
        }
        tmem_st16(s_addr + c * kKC, pk);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.p_ready[c]));
      };
      {
        uint32_t ra[32], rb[32];
        tmem_ld32(s_addr, ra);
        for (int c = 0; c < nC; c += 2) {
          tmem_ld_wait();
          if (c + 1 < nC) tmem_ld32(s_addr + (c + 1) * kKC, rb);
          exp_chunk(ra, c);
          if (c + 1 < nC) {
            tmem_ld_wait();
            if (c + 2 < nC) tmem_ld32(s_addr + (c + 2) * kKC, ra);
            exp_chunk(rb, c + 1);
          }
        }
      }
      if (te == 0) TRACE(0, 14);
      S.xl[wg][row] = sum;
      named_bar_sync(1, kSoftThreads);
      sum += S.xl[wg ^ 1][row];
      if (wg == 0 && lse != nullptr && w.q0 + row < Nq)      // (under the wait for the last P V)
        lse[((int64_t)w.b * H + w.h) * Nq + w.q0 + row] = 0.6931471805599453f * (mx + log2f(sum));
      // ---- epilogue: O / l -> bf16 -> this item's Q buffer (half wg) -> store thread
      if (te == 0) TRACE(0, 15);
      mbar_wait(smem_u32(&S.o_full), qi & 1);
      tc_fence_after();
      if (te == 0) TRACE(0, 16);
      const float rinv = __frcp_rn(sum);
      {
        uint32_t ra[32], rb[32];
        tmem_ld32(tmem + lane_base + kColO + wg * 64, ra);
        tmem_ld32(tmem + lane_base + kColO + wg * 64 + 32, rb);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.o_empty));     // O is in registers: the next item's P V may start
        uint8_t* orow = &S.q[qi & 1][wg * kQHalf + row * 128];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t wa[4], wb[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            wa[j] = pack_bf16(__uint_as_float(ra[g * 8 + 2 * j]) * rinv, __uint_as_float(ra[g * 8 + 2 * j + 1]) * rinv);
            wb[j] = pack_bf16(__uint_as_float(rb[g * 8 + 2 * j]) * rinv, __uint_as_float(rb[g * 8 + 2 * j + 1]) * rinv);
          }
          *reinterpret_cast<uint4*>(orow + ((g ^ (row & 7)) << 4)) = make_uint4(wa[0], wa[1], wa[2], wa[3]);
          *reinterpret_cast<uint4*>(orow + (((4 + g) ^ (row & 7)) << 4)) = make_uint4(wb[0], wb[1], wb[2], wb[3]);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&S.o_staged));
      if (te == 0) TRACE(0, 17);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

}  // namespace
}  // namespace yat

// ------------------------------------------------------------------ host side
YAT_DEFINE_TRACE_API(yat_xattn)

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
  if ((rc = make_tmap_heads_bf16(&tk, k, d, H, L, B, (uint64_t)ldkv, kKR)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tv, v, d, H, L, B, (uint64_t)ldkv, kKR)) != YAT_OK) return rc;
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
