// Cross-attention of the SANA / PixArt blocks (attn2), backward, on the 5th-gen tensor cores (tcgen05 + TMEM + TMA).
//
// With s_ij = scale q_i.k_j + bias_j, p_ij = exp(s_ij - lse_i) (lse from the forward, yat_xattn_fwd) and D_i = dO_i . O_i:
//   dP_ij = dO_i . v_j        dS_ij = p_ij (dP_ij - D_i)
//   dQ_i = scale sum_j dS_ij k_j        dK_j = scale sum_i dS_ij q_i        dV_j = sum_i p_ij dO_i
// (the backward of F.scaled_dot_product_attention inside diffusers' AttnProcessor2_0, reached from
// utils/patch_sana_attention_layers.py:100-106).  Three kernels:
//   prep   D[b, h, i] = sum_c dO . O                                   (coalesced slab loads, fixed-order sums; HBM-bound)
//   dq     item = (batch, head, 128-query tile): S_c = Q K_c^T and dP_c = dO V_c^T per 128-key chunk, p kept in registers,
//          dS_c written back over dP_c by the threads that own the rows (bf16 pairs), dQ += dS_c K_c with its A operand read
//          from tensor memory - the same feedback as the forward's P V
//   dkv    item = (batch, head, 128-key block), K and V blocks resident, looping over the query tiles: S^T = K Q_t^T and
//          dP^T = V dO_t^T (keys along the lanes), P^T and dS^T written back over them, dV += P^T dO_t, dK += dS^T Q_t
// No probability ever touches shared memory; every reduction over keys or queries is an MMA.
#include "common.cuh"
#include "sm100.cuh"

namespace yat {
namespace {
using namespace sm100;

constexpr int kQT = 128;                     // queries per tile
constexpr int kKC = 64;                      // keys per sub-chunk (dq kernel)
constexpr int kMaxKeys = 384;
constexpr int kTHalf = kQT * 128;            // [128 rows][64 cols] bf16
constexpr int kTBytes = 2 * kTHalf;          // 32 KB
constexpr int kSoftWarps = 8, kSoftThreads = 32 * kSoftWarps;
constexpr int kStoreWarp = 2 + kSoftWarps;
constexpr int kThreads = 32 * (kStoreWarp + 1);
constexpr int kTmemCols = 512;
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ------------------------------------------------------------------------------------------------ prep: D = rowsum(dO * O)
// Block = 8 token rows: coalesced 128-bit loads over the whole [8, H*d] slab (one fp32 partial per 8 elements into shared
// memory), then one thread per (row, head) adds that head's d / 8 partials in a fixed order.
constexpr int kPrepRows = 8;
__global__ void __launch_bounds__(256) xattn_prep_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                                                         float* __restrict__ D, int64_t rows, int H, int d, int Nq) {
  extern __shared__ float part[];                                         // [kPrepRows][H * d / 8]
  const int vec_per_row = H * d / 8, vec_per_head = d / 8;
  const int64_t r0 = (int64_t)blockIdx.x * kPrepRows;
  const int nrows = (int)(rows - r0 < kPrepRows ? rows - r0 : kPrepRows);
  const uint4* po = reinterpret_cast<const uint4*>(o + r0 * H * d);
  const uint4* pd = reinterpret_cast<const uint4*>(d_o + r0 * H * d);
  for (int i = threadIdx.x; i < nrows * vec_per_row; i += 256) {
    const uint4 a = ld_stream(po + i), b = ld_stream(pd + i);
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) acc += bf16_lo(aw[j]) * bf16_lo(bw[j]) + bf16_hi(aw[j]) * bf16_hi(bw[j]);
    part[i] = acc;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nrows * H; i += 256) {
    const int rr = i / H, h = i - rr * H;
    const float* p = part + rr * vec_per_row + h * vec_per_head;
    float acc = 0.f;
    for (int j = 0; j < vec_per_head; ++j) acc += p[j];
    const int64_t r = r0 + rr;
    D[((r / Nq) * H + h) * Nq + (r % Nq)] = acc;                          // [B, H, Nq] like lse
  }
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

// ------------------------------------------------------------------------------------------------ dq kernel
// 128-key ring chunks (N = 128 MMAs: a 64-key MMA with both operands in shared memory is bound by its operand reads), single
// TMEM buffers, two-phase hand-shake: phase 1 reads S_c and keeps p in registers, phase 2 reads dP_c and writes dS_c over the
// dP columns; the in-order tensor pipe overlaps everything:  S_0 dP_0 | [p_c] S_{c+1} | [dS_c] dQ_c dP_{c+1} | ...
constexpr int kKR = 128;                      // keys per ring chunk
constexpr int kRHalf = kKR * 128;             // [128 keys][64 cols] bf16
constexpr int kRBytes = 2 * kRHalf;           // 32 KB
constexpr int kDqStages = 3;                  // V_c K_c V_{c+1}: V_c is free after dP_c (early), K_c after dQ_c
constexpr uint32_t kColS = 0, kColDP = 128, kColDQ = 256;

struct DqSmem {
  uint8_t qd[2][2][kTBytes];                 // [item buffer][Q | dO]; the dQ tile leaves from the Q half
  uint8_t kv[kDqStages][kRBytes];
  float bias2[kMaxKeys];
  uint64_t qd_full[2], kv_full[kDqStages], kv_empty[kDqStages], s_full, dp_full, p1_done, ds_ready;
  uint64_t dq_full, dq_empty, dq_staged;
  uint32_t tmem_base;
};
constexpr int kDqSmemBytes = (int)sizeof(DqSmem) + 1024;
static_assert(kDqSmemBytes <= 232448, "xattn dq kernel exceeds the 227 KB shared-memory limit");

constexpr uint32_t kIdescSd128 = make_idesc_bf16(128, 128, 0, 0);   // Q K_c^T, dO V_c^T : A, B K-major
constexpr uint32_t kIdescSd64 = make_idesc_bf16(128, 64, 0, 0);     // last chunk with <= 64 keys

__global__ void __launch_bounds__(kThreads, 1)
xattn_dq_sm100_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_do,
                      const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                      const __grid_constant__ CUtensorMap tm_dq, const float* __restrict__ bias,
                      const float* __restrict__ lse, const float* __restrict__ D, int B, int H, int Nq, int L, int d,
                      float scale) {
  extern __shared__ uint8_t smem_raw[];
  DqSmem& S = *reinterpret_cast<DqSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_qt = (Nq + kQT - 1) / kQT;
  const int n_items = B * H * n_qt;
  const int nC = (L + kKC - 1) / kKC;          // 64-key sub-chunks
  const int nR = (L + kKR - 1) / kKR;          // 128-key ring chunks
  const int nK = (d + 15) / 16;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) mbar_init(smem_u32(&S.qd_full[i]), 1);
    for (int s = 0; s < kDqStages; ++s) {
      mbar_init(smem_u32(&S.kv_full[s]), 1);
      mbar_init(smem_u32(&S.kv_empty[s]), 1);
    }
    mbar_init(smem_u32(&S.s_full), 1);
    mbar_init(smem_u32(&S.dp_full), 1);
    mbar_init(smem_u32(&S.p1_done), kSoftWarps);
    mbar_init(smem_u32(&S.ds_ready), kSoftWarps);
    mbar_init(smem_u32(&S.dq_full), 1);
    mbar_init(smem_u32(&S.dq_empty), kSoftWarps);
    mbar_init(smem_u32(&S.dq_staged), kSoftWarps);
    fence_barrier_init();
    tma_prefetch_desc(&tm_q); tma_prefetch_desc(&tm_do); tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v);
    tma_prefetch_desc(&tm_dq);
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(&S.tmem_base), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (S.tmem_base != 0) __trap();
  constexpr uint32_t tmem = 0;

  if (warp == 0) {
    // ======================= TMA producer: the key ring (V_c, K_c per chunk) =======================
    if (lane == 0) {
      uint32_t it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const Item w = decode(item, n_qt, H);
        for (int n = 0; n < 2 * nR; ++n, ++it) {
          const CUtensorMap* tm = (n & 1) ? &tm_k : &tm_v;
          const uint32_t s = it % kDqStages, ph = (it / kDqStages) & 1;
          mbar_wait(smem_u32(&S.kv_empty[s]), ph ^ 1);
          const uint32_t full = smem_u32(&S.kv_full[s]);
          mbar_arrive_expect_tx(full, kRBytes);
          tma_load_4d(smem_u32(&S.kv[s][0]), tm, full, 0, w.h, (n >> 1) * kKR, w.b);
          tma_load_4d(smem_u32(&S.kv[s][kRHalf]), tm, full, 64, w.h, (n >> 1) * kKR, w.b);
        }
      }
    }
  } else if (warp == 1) {
    // ======================= tcgen05 issuer =======================
    if (lane == 0) {
      uint32_t it0 = 0, qi = 0, cc = 0;            // cc: chunks handed to the softmax warps so far (parity of their barriers)
      const uint32_t idesc_dq = make_idesc_bf16(128, (d + 15) & ~15, 0, 1);
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi, it0 += 2 * nR) {
        const uint32_t buf = qi & 1;
        mbar_wait(smem_u32(&S.qd_full[buf]), (qi >> 1) & 1);
        tc_fence_after();
        const uint64_t qa0 = make_smem_desc(smem_u32(&S.qd[buf][0][0]), 16, 1024, kLayoutSw128);
        const uint64_t da0 = make_smem_desc(smem_u32(&S.qd[buf][1][0]), 16, 1024, kLayoutSw128);
        auto idesc_of = [&](int c) { return (nC - 2 * c >= 2) ? kIdescSd128 : kIdescSd64; };
        auto issue_s = [&](int c) {                 // S_c = Q K_c^T
          const uint32_t ik = it0 + 2 * c + 1, sk = ik % kDqStages;
          mbar_wait(smem_u32(&S.kv_full[sk]), (ik / kDqStages) & 1);
          tc_fence_after();
          const uint64_t kb0 = make_smem_desc(smem_u32(&S.kv[sk][0]), 16, 1024, kLayoutSw128);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            if (ks < nK)
              umma_bf16(tmem + kColS, qa0 + (((ks >> 2) * kTHalf + (ks & 3) * 32) >> 4),
                        kb0 + (((ks >> 2) * kRHalf + (ks & 3) * 32) >> 4), idesc_of(c), ks != 0);
          tc_commit(smem_u32(&S.s_full));
        };
        auto issue_dp = [&](int c) {                // dP_c = dO V_c^T
          const uint32_t iv = it0 + 2 * c, sv = iv % kDqStages;
          mbar_wait(smem_u32(&S.kv_full[sv]), (iv / kDqStages) & 1);
          tc_fence_after();
          const uint64_t vb0 = make_smem_desc(smem_u32(&S.kv[sv][0]), 16, 1024, kLayoutSw128);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            if (ks < nK)
              umma_bf16(tmem + kColDP, da0 + (((ks >> 2) * kTHalf + (ks & 3) * 32) >> 4),
                        vb0 + (((ks >> 2) * kRHalf + (ks & 3) * 32) >> 4), idesc_of(c), ks != 0);
          tc_commit(smem_u32(&S.kv_empty[sv]));        // V_c is done
          tc_commit(smem_u32(&S.dp_full));
        };
        issue_s(0);
        issue_dp(0);
        for (int c = 0; c < nR; ++c, ++cc) {
          mbar_wait(smem_u32(&S.p1_done), cc & 1);     // every softmax warp has read S_c
          if (c + 1 < nR) issue_s(c + 1);
          mbar_wait(smem_u32(&S.ds_ready), cc & 1);    // dS_c is in the dP columns
          if (c == 0) mbar_wait(smem_u32(&S.dq_empty), (qi & 1) ^ 1);
          tc_fence_after();
          // dQ += dS_c K_c: A = dS_c in tensor memory (16 keys of K step ks as 8 packed columns at
          // 64 (ks >> 2) + 32 ((ks >> 1) & 1) + 8 (ks & 1) of the dP region), B = K_c MN-major (K step of 16 keys = 2 KB)
          const uint32_t ik = it0 + 2 * c + 1, sk = ik % kDqStages;
          const uint64_t kb0 = make_smem_desc(smem_u32(&S.kv[sk][0]), kRHalf, 1024, kLayoutSw128);
          const int nks = (nC - 2 * c >= 2) ? 8 : 4;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            if (ks < nks)
              umma_bf16_ts(tmem + kColDQ, tmem + kColDP + 64 * (ks >> 2) + 32 * ((ks >> 1) & 1) + 8 * (ks & 1),
                           kb0 + ((ks * 2048) >> 4), idesc_dq, (c | ks) != 0);
          tc_commit(smem_u32(&S.kv_empty[sk]));
          if (c == nR - 1) tc_commit(smem_u32(&S.dq_full));
          if (c + 1 < nR) issue_dp(c + 1);             // behind dQ_c: the dP columns are free
        }
      }
    }
  } else if (warp == kStoreWarp) {
    // ======================= Q / dO loads + dQ stores (one thread) =======================
    if (lane == 0) {
      auto load_qd = [&](int item, uint32_t buf) {
        const Item w = decode(item, n_qt, H);
        const uint32_t f = smem_u32(&S.qd_full[buf]);
        mbar_arrive_expect_tx(f, 2 * kTBytes);
        tma_load_4d(smem_u32(&S.qd[buf][0][0]), &tm_q, f, 0, w.h, w.q0, w.b);
        tma_load_4d(smem_u32(&S.qd[buf][0][kTHalf]), &tm_q, f, 64, w.h, w.q0, w.b);
        tma_load_4d(smem_u32(&S.qd[buf][1][0]), &tm_do, f, 0, w.h, w.q0, w.b);
        tma_load_4d(smem_u32(&S.qd[buf][1][kTHalf]), &tm_do, f, 64, w.h, w.q0, w.b);
      };
      if ((int)blockIdx.x < n_items) load_qd(blockIdx.x, 0);
      if ((int)(blockIdx.x + gridDim.x) < n_items) load_qd(blockIdx.x + gridDim.x, 1);
      uint32_t qi = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi) {
        const Item w = decode(item, n_qt, H);
        const uint32_t buf = qi & 1;
        mbar_wait(smem_u32(&S.dq_staged), qi & 1);
        tma_store_4d(&tm_dq, smem_u32(&S.qd[buf][0][0]), 0, w.h, w.q0, w.b);
        if (d > 64) tma_store_4d(&tm_dq, smem_u32(&S.qd[buf][0][kTHalf]), 64, w.h, w.q0, w.b);
        tma_store_commit();
        tma_store_wait_read<0>();
        if (item + 2 * (int)gridDim.x < n_items) load_qd(item + 2 * gridDim.x, buf);
      }
      tma_store_wait_all<0>();
    }
  } else {
    // ======================= p / dS + epilogue warps =======================
    const int te = threadIdx.x - 64;              // 0..255
    const int wg = te >> 7;                       // which 64 of a chunk's 128 keys / which half of dQ
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const f2 cs2 = bc2(scale * kLog2e);
    uint32_t qi = 0, cc = 0;
    // per-item scalars and the sample's bias row are fetched one item ahead (registers): no global latency between two items
    auto fetch = [&](int item, float& l2, float& dn, float& b0, float& b1) {
      const Item w = decode(item, n_qt, H);
      const bool valid = w.q0 + row < Nq;
      const int64_t vi = ((int64_t)w.b * H + w.h) * Nq + w.q0 + row;
      l2 = valid ? __ldg(lse + vi) * kLog2e : 0.f;
      dn = valid ? -__ldg(D + vi) : 0.f;
      b0 = b1 = 0.f;
      if (bias != nullptr) {
        if (te < L) b0 = __ldg(bias + (int64_t)w.b * L + te);
        if (te + kSoftThreads < L) b1 = __ldg(bias + (int64_t)w.b * L + te + kSoftThreads);
      }
    };
    auto put_bias = [&](float b0, float b1) {
      if (te < nR * kKR) S.bias2[te] = te < L ? b0 * kLog2e : -INFINITY;
      if (te + kSoftThreads < nR * kKR) S.bias2[te + kSoftThreads] = te + kSoftThreads < L ? b1 * kLog2e : -INFINITY;
    };
    float n_l2 = 0.f, n_dn = 0.f, n_b0 = 0.f, n_b1 = 0.f;
    if ((int)blockIdx.x < n_items) {
      fetch(blockIdx.x, n_l2, n_dn, n_b0, n_b1);
      put_bias(n_b0, n_b1);
      named_bar_sync(1, kSoftThreads);
    }
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++qi) {
      const float lse2 = n_l2;
      const f2 nd2 = bc2(n_dn);
      const bool has_next = item + (int)gridDim.x < n_items;
      if (has_next) fetch(item + gridDim.x, n_l2, n_dn, n_b0, n_b1);
#ifdef YAT_BWD_PROBE
      long long w_s = 0, w_dp = 0, w_s0 = 0;
      if (te == 0) TRACE(0, 10);
#define XW(acc, ...) do { const long long t0_ = clock64(); __VA_ARGS__; acc += clock64() - t0_; } while (0)
#else
#define XW(acc, ...) do { __VA_ARGS__; } while (0)
#endif
      for (int c = 0; c < nR; ++c, ++cc) {
        // phase 1: p = 2^(s c + bias2 - lse2), kept in registers
        if (c == 0) XW(w_s0, mbar_wait(smem_u32(&S.s_full), cc & 1)); else XW(w_s, mbar_wait(smem_u32(&S.s_full), cc & 1));
        tc_fence_after();
        f2 pr[2][16];
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint32_t rs[32];
          const uint32_t col = wg * 64 + hh * 32;
          tmem_ld32(tmem + lane_base + kColS + col, rs);
          tmem_ld_wait();
          const float4* b4 = reinterpret_cast<const float4*>(&S.bias2[c * kKR + col]);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 bb = b4[i];
            float a0, a1, a2, a3;
            un2(fma2(mk2(__uint_as_float(rs[4 * i]), __uint_as_float(rs[4 * i + 1])), cs2, mk2(bb.x - lse2, bb.y - lse2)), a0, a1);
            un2(fma2(mk2(__uint_as_float(rs[4 * i + 2]), __uint_as_float(rs[4 * i + 3])), cs2, mk2(bb.z - lse2, bb.w - lse2)), a2, a3);
            pr[hh][2 * i] = mk2(ex2(a0), ex2(a1));
            pr[hh][2 * i + 1] = mk2(ex2(a2), ex2(a3));
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.p1_done));
        // phase 2: dS = p (dP - D) -> bf16 pairs over the first 16 of each 32 dP columns
        XW(w_dp, mbar_wait(smem_u32(&S.dp_full), cc & 1));
        tc_fence_after();
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint32_t rp[32];
          const uint32_t col = wg * 64 + hh * 32;
          tmem_ld32(tmem + lane_base + kColDP + col, rp);
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            pk[2 * i] = pack2(mul2(pr[hh][2 * i], add2(mk2(__uint_as_float(rp[4 * i]), __uint_as_float(rp[4 * i + 1])), nd2)));
            pk[2 * i + 1] = pack2(mul2(pr[hh][2 * i + 1], add2(mk2(__uint_as_float(rp[4 * i + 2]), __uint_as_float(rp[4 * i + 3])), nd2)));
          }
          tmem_st16(tmem + lane_base + kColDP + col, pk);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.ds_ready));
      }
      named_bar_sync(1, kSoftThreads);                  // every warp has finished reading this item's bias row
      if (has_next) put_bias(n_b0, n_b1);
#ifdef YAT_BWD_PROBE
      if (te == 0) { TRACE(0, 1000 + (int)(w_s >> 4)); TRACE(0, 2000 + (int)(w_dp >> 4)); TRACE(0, 3000 + (int)(w_s0 >> 4)); TRACE(0, 15); }
#endif
      // ---- epilogue: scale dQ -> bf16 -> the item's Q buffer -> store thread
      mbar_wait(smem_u32(&S.dq_full), qi & 1);
      tc_fence_after();
      if (te == 0) TRACE(0, 16);
      {
        uint32_t ra[32], rb[32];
        tmem_ld32(tmem + lane_base + kColDQ + wg * 64, ra);
        tmem_ld32(tmem + lane_base + kColDQ + wg * 64 + 32, rb);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.dq_empty));
        uint8_t* orow = &S.qd[qi & 1][0][wg * kTHalf + row * 128];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t wa[4], wb[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            wa[i] = pack_bf16(__uint_as_float(ra[g * 8 + 2 * i]) * scale, __uint_as_float(ra[g * 8 + 2 * i + 1]) * scale);
            wb[i] = pack_bf16(__uint_as_float(rb[g * 8 + 2 * i]) * scale, __uint_as_float(rb[g * 8 + 2 * i + 1]) * scale);
          }
          *reinterpret_cast<uint4*>(orow + ((g ^ (row & 7)) << 4)) = make_uint4(wa[0], wa[1], wa[2], wa[3]);
          *reinterpret_cast<uint4*>(orow + (((4 + g) ^ (row & 7)) << 4)) = make_uint4(wb[0], wb[1], wb[2], wb[3]);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&S.dq_staged));
      if (te == 0) TRACE(0, 17);
      named_bar_sync(1, kSoftThreads);                  // the next item's bias row is in place
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

// ------------------------------------------------------------------------------------------------ dkv kernel
constexpr int kKB = 128;                      // keys per block (the lanes of S^T, dP^T, dV, dK)
constexpr int kQdStages = 2;
constexpr uint32_t kColST = 0, kColDPT = 128, kColDV = 256, kColDK = 384;

struct DkvSmem {
  uint8_t kvb[2][kTBytes];                    // K block | V block ([2 halves][128 keys][64]); dK | dV leave from here
  uint8_t qd[kQdStages][2][kTBytes];          // ring of query tiles: Q_t | dO_t
  float lq[kQdStages][kQT], dd[kQdStages][kQT];   // per query of the tile: lse * log2(e), D
  uint64_t kvb_full, kvb_empty, qd_full[kQdStages], qd_empty[kQdStages], st_full, dpt_full, p_ready, ds_ready;
  uint64_t dkv_full, dkv_empty, dkv_staged;
  uint32_t tmem_base;
};
constexpr int kDkvSmemBytes = (int)sizeof(DkvSmem) + 1024;
static_assert(kDkvSmemBytes <= 232448, "xattn dkv kernel exceeds the 227 KB shared-memory limit");

constexpr uint32_t kIdescTT = make_idesc_bf16(128, 128, 0, 0);   // K Q_t^T, V dO_t^T : A, B K-major

struct KItem { int b, h, kb; };
__device__ __forceinline__ KItem kdecode(int item, int n_kb, int H) {
  KItem r;
  r.kb = item % n_kb;
  const int bh = item / n_kb;
  r.h = bh % H;
  r.b = bh / H;
  return r;
}

__global__ void __launch_bounds__(kThreads, 1)
xattn_dkv_sm100_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_do,
                       const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                       const __grid_constant__ CUtensorMap tm_dk, const __grid_constant__ CUtensorMap tm_dv,
                       const float* __restrict__ bias, const float* __restrict__ lse, const float* __restrict__ D, int B, int H,
                       int Nq, int L, int d, float scale) {
  extern __shared__ uint8_t smem_raw[];
  DkvSmem& S = *reinterpret_cast<DkvSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_kb = (L + kKB - 1) / kKB, nT = (Nq + kQT - 1) / kQT;
  const int n_items = B * H * n_kb;
  const int nK = (d + 15) / 16;

  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&S.kvb_full), 1);
    mbar_init(smem_u32(&S.kvb_empty), 1);
    for (int s = 0; s < kQdStages; ++s) {
      mbar_init(smem_u32(&S.qd_full[s]), 1);
      mbar_init(smem_u32(&S.qd_empty[s]), 1);
    }
    mbar_init(smem_u32(&S.st_full), 1);
    mbar_init(smem_u32(&S.dpt_full), 1);
    mbar_init(smem_u32(&S.p_ready), kSoftWarps);
    mbar_init(smem_u32(&S.ds_ready), kSoftWarps);
    mbar_init(smem_u32(&S.dkv_full), 1);
    mbar_init(smem_u32(&S.dkv_empty), kSoftWarps);
    mbar_init(smem_u32(&S.dkv_staged), kSoftWarps);
    fence_barrier_init();
    tma_prefetch_desc(&tm_q); tma_prefetch_desc(&tm_do); tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v);
    tma_prefetch_desc(&tm_dk); tma_prefetch_desc(&tm_dv);
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(&S.tmem_base), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (S.tmem_base != 0) __trap();
  constexpr uint32_t tmem = 0;

  if (warp == 0) {
    // ======================= TMA producer =======================
    if (lane == 0) {
      uint32_t tt = 0, ii = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++ii) {
        const KItem w = kdecode(item, n_kb, H);
        mbar_wait(smem_u32(&S.kvb_empty), (ii & 1) ^ 1);
        const uint32_t f = smem_u32(&S.kvb_full);
        mbar_arrive_expect_tx(f, 2 * kTBytes);
        tma_load_4d(smem_u32(&S.kvb[0][0]), &tm_k, f, 0, w.h, w.kb * kKB, w.b);
        tma_load_4d(smem_u32(&S.kvb[0][kTHalf]), &tm_k, f, 64, w.h, w.kb * kKB, w.b);
        tma_load_4d(smem_u32(&S.kvb[1][0]), &tm_v, f, 0, w.h, w.kb * kKB, w.b);
        tma_load_4d(smem_u32(&S.kvb[1][kTHalf]), &tm_v, f, 64, w.h, w.kb * kKB, w.b);
        for (int t = 0; t < nT; ++t, ++tt) {
          const uint32_t s = tt % kQdStages, ph = (tt / kQdStages) & 1;
          mbar_wait(smem_u32(&S.qd_empty[s]), ph ^ 1);
          const uint32_t qf = smem_u32(&S.qd_full[s]);
          mbar_arrive_expect_tx(qf, 2 * kTBytes);
          tma_load_4d(smem_u32(&S.qd[s][0][0]), &tm_q, qf, 0, w.h, t * kQT, w.b);
          tma_load_4d(smem_u32(&S.qd[s][0][kTHalf]), &tm_q, qf, 64, w.h, t * kQT, w.b);
          tma_load_4d(smem_u32(&S.qd[s][1][0]), &tm_do, qf, 0, w.h, t * kQT, w.b);
          tma_load_4d(smem_u32(&S.qd[s][1][kTHalf]), &tm_do, qf, 64, w.h, t * kQT, w.b);
        }
      }
    }
  } else if (warp == 1) {
    // ======================= tcgen05 issuer =======================
    if (lane == 0) {
      uint32_t tt = 0, ii = 0;
      const uint32_t idesc_acc = make_idesc_bf16(128, (d + 15) & ~15, 0, 1);
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++ii) {
        mbar_wait(smem_u32(&S.kvb_full), ii & 1);
        tc_fence_after();
        const uint64_t ka0 = make_smem_desc(smem_u32(&S.kvb[0][0]), 16, 1024, kLayoutSw128);
        const uint64_t va0 = make_smem_desc(smem_u32(&S.kvb[1][0]), 16, 1024, kLayoutSw128);
        // Two-phase hand-shake with the softmax warps (phase 1 reads S^T and writes P^T over it, phase 2 reads dP^T and writes
        // dS^T over it), single TMEM buffers, everything overlapped through the in-order tensor pipe:
        //   S^T_0 dP^T_0 | [P^T_t] dV_t  S^T_{t+1} | [dS^T_t] dK_t  dP^T_{t+1} | ...
        auto issue_st = [&](uint32_t n) {          // S^T = K Q_t^T (keys along the lanes, the tile's queries along the columns)
          const uint32_t s = n % kQdStages;
          mbar_wait(smem_u32(&S.qd_full[s]), (n / kQdStages) & 1);
          tc_fence_after();
          const uint64_t qb0 = make_smem_desc(smem_u32(&S.qd[s][0][0]), 16, 1024, kLayoutSw128);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            if (ks < nK) {
              const uint32_t off = ((ks >> 2) * kTHalf + (ks & 3) * 32) >> 4;
              umma_bf16(tmem + kColST, ka0 + off, qb0 + off, kIdescTT, ks != 0);
            }
          tc_commit(smem_u32(&S.st_full));
        };
        auto issue_dpt = [&](uint32_t n) {         // dP^T = V dO_t^T (the tile's qd_full has been observed by issue_st)
          const uint32_t s = n % kQdStages;
          const uint64_t db0 = make_smem_desc(smem_u32(&S.qd[s][1][0]), 16, 1024, kLayoutSw128);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            if (ks < nK) {
              const uint32_t off = ((ks >> 2) * kTHalf + (ks & 3) * 32) >> 4;
              umma_bf16(tmem + kColDPT, va0 + off, db0 + off, kIdescTT, ks != 0);
            }
          tc_commit(smem_u32(&S.dpt_full));
        };
        issue_st(tt);
        issue_dpt(tt);
        for (int t = 0; t < nT; ++t, ++tt) {
          const uint32_t s = tt % kQdStages;
          // dV += P^T dO_t, dK += dS^T Q_t: A from tensor memory (16 queries of K step ks as 8 packed columns at
          // 64 (ks >> 2) + 32 ((ks >> 1) & 1) + 8 (ks & 1)), B = the tile MN-major (K step of 16 query rows = 2 KB)
          const uint64_t dob = make_smem_desc(smem_u32(&S.qd[s][1][0]), kTHalf, 1024, kLayoutSw128);
          const uint64_t qmb = make_smem_desc(smem_u32(&S.qd[s][0][0]), kTHalf, 1024, kLayoutSw128);
          mbar_wait(smem_u32(&S.p_ready), tt & 1);
          if (t == 0) mbar_wait(smem_u32(&S.dkv_empty), (ii & 1) ^ 1);    // the previous item's dV / dK have left TMEM
          tc_fence_after();
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint32_t acol = 64 * (ks >> 2) + 32 * ((ks >> 1) & 1) + 8 * (ks & 1);
            umma_bf16_ts(tmem + kColDV, tmem + kColST + acol, dob + ((ks * 2048) >> 4), idesc_acc, (t | ks) != 0);
          }
          if (t + 1 < nT) issue_st(tt + 1);         // behind dV_t in the pipe: the S^T columns are free
          mbar_wait(smem_u32(&S.ds_ready), tt & 1);
          tc_fence_after();
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint32_t acol = 64 * (ks >> 2) + 32 * ((ks >> 1) & 1) + 8 * (ks & 1);
            umma_bf16_ts(tmem + kColDK, tmem + kColDPT + acol, qmb + ((ks * 2048) >> 4), idesc_acc, (t | ks) != 0);
          }
          tc_commit(smem_u32(&S.qd_empty[s]));
          if (t + 1 < nT) issue_dpt(tt + 1);        // behind dK_t: the dP^T columns are free
          if (t == nT - 1) tc_commit(smem_u32(&S.dkv_full));
        }
      }
    }
  } else if (warp == kStoreWarp) {
    // ======================= dK / dV stores =======================
    if (lane == 0) {
      uint32_t ii = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++ii) {
        const KItem w = kdecode(item, n_kb, H);
        mbar_wait(smem_u32(&S.dkv_staged), ii & 1);
        tma_store_4d(&tm_dk, smem_u32(&S.kvb[0][0]), 0, w.h, w.kb * kKB, w.b);
        tma_store_4d(&tm_dv, smem_u32(&S.kvb[1][0]), 0, w.h, w.kb * kKB, w.b);
        if (d > 64) {
          tma_store_4d(&tm_dk, smem_u32(&S.kvb[0][kTHalf]), 64, w.h, w.kb * kKB, w.b);
          tma_store_4d(&tm_dv, smem_u32(&S.kvb[1][kTHalf]), 64, w.h, w.kb * kKB, w.b);
        }
        tma_store_commit();
        tma_store_wait_read<0>();
        mbar_arrive(smem_u32(&S.kvb_empty));        // the next item's K / V blocks may land
      }
      tma_store_wait_all<0>();
    }
  } else {
    // ======================= P^T / dS^T + epilogue warps =======================
    const int te = threadIdx.x - 64;              // 0..255
    const int wg = te >> 7;                       // which 64 of the tile's 128 queries / which half of dV, dK
    const int quad = warp & 3;
    const int row = quad * 32 + lane;             // key row of the block (TMEM lane)
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const f2 cs2 = bc2(scale * kLog2e);
    uint32_t tt = 0, ii = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++ii) {
      const KItem w = kdecode(item, n_kb, H);
      const int key = w.kb * kKB + row;
      const float bk = key < L ? (bias ? __ldg(bias + (int64_t)w.b * L + key) * kLog2e : 0.f) : -INFINITY;
      const float* lrow = lse + ((int64_t)w.b * H + w.h) * Nq;
      const float* drow = D + ((int64_t)w.b * H + w.h) * Nq;
      // per-query vectors of tile 0 (the vectors of tile t + 1 are fetched while tile t is processed)
      {
        const int qq = te & 127;
        const float x = qq < Nq ? (te < 128 ? __ldg(lrow + qq) * kLog2e : __ldg(drow + qq)) : 0.f;
        (te < 128 ? S.lq : S.dd)[tt % kQdStages][qq] = x;
        named_bar_sync(1, kSoftThreads);
      }
      for (int t = 0; t < nT; ++t, ++tt) {
        const uint32_t s = tt % kQdStages;
        float nxt = 0.f;
        if (t + 1 < nT) {
          const int qq = (t + 1) * kQT + (te & 127);
          if (qq < Nq) nxt = te < 128 ? __ldg(lrow + qq) * kLog2e : __ldg(drow + qq);
        }
        // phase 1: S^T -> p (kept in registers), P^T (bf16 pairs) over the first 16 of each 32 score columns
        mbar_wait(smem_u32(&S.st_full), tt & 1);
        tc_fence_after();
        f2 pr[2][16];
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint32_t rs[32];
          const uint32_t col = wg * 64 + hh * 32;
          tmem_ld32(tmem + lane_base + kColST + col, rs);
          tmem_ld_wait();
          const float4* l4 = reinterpret_cast<const float4*>(&S.lq[s][col]);
          uint32_t pp[16];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 ll = l4[i];
            float a0, a1, a2, a3;
            un2(fma2(mk2(__uint_as_float(rs[4 * i]), __uint_as_float(rs[4 * i + 1])), cs2, mk2(bk - ll.x, bk - ll.y)), a0, a1);
            un2(fma2(mk2(__uint_as_float(rs[4 * i + 2]), __uint_as_float(rs[4 * i + 3])), cs2, mk2(bk - ll.z, bk - ll.w)), a2, a3);
            pr[hh][2 * i] = mk2(ex2(a0), ex2(a1));
            pr[hh][2 * i + 1] = mk2(ex2(a2), ex2(a3));
            pp[2 * i] = pack2(pr[hh][2 * i]);
            pp[2 * i + 1] = pack2(pr[hh][2 * i + 1]);
          }
          tmem_st16(tmem + lane_base + kColST + col, pp);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.p_ready));
        // phase 2: dP^T -> dS^T = p (dP^T - D) over the dP^T columns
        mbar_wait(smem_u32(&S.dpt_full), tt & 1);
        tc_fence_after();
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint32_t rp[32];
          const uint32_t col = wg * 64 + hh * 32;
          tmem_ld32(tmem + lane_base + kColDPT + col, rp);
          tmem_ld_wait();
          const float4* d4 = reinterpret_cast<const float4*>(&S.dd[s][col]);
          uint32_t ps[16];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 dv4 = d4[i];
            ps[2 * i] = pack2(mul2(pr[hh][2 * i], add2(mk2(__uint_as_float(rp[4 * i]), __uint_as_float(rp[4 * i + 1])), mk2(-dv4.x, -dv4.y))));
            ps[2 * i + 1] = pack2(mul2(pr[hh][2 * i + 1], add2(mk2(__uint_as_float(rp[4 * i + 2]), __uint_as_float(rp[4 * i + 3])), mk2(-dv4.z, -dv4.w))));
          }
          tmem_st16(tmem + lane_base + kColDPT + col, ps);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.ds_ready));
        if (t + 1 < nT) {                       // the other stage's vectors were last read two tiles ago
          (te < 128 ? S.lq : S.dd)[(tt + 1) % kQdStages][te & 127] = nxt;
          named_bar_sync(1, kSoftThreads);
        }
      }
      // ---- epilogue: dV, scale dK -> bf16 -> the K / V block buffers -> store warp
      mbar_wait(smem_u32(&S.dkv_full), ii & 1);
      tc_fence_after();
#pragma unroll
      for (int which = 0; which < 2; ++which) {   // 0: dK (scaled) into kvb[0]; 1: dV into kvb[1]
        uint32_t ra[32], rb[32];
        const uint32_t col = (which ? kColDV : kColDK) + wg * 64;
        tmem_ld32(tmem + lane_base + col, ra);
        tmem_ld32(tmem + lane_base + col + 32, rb);
        tmem_ld_wait();
        const float f = which ? 1.f : scale;
        uint8_t* orow = &S.kvb[which][wg * kTHalf + row * 128];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t wa[4], wb[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            wa[i] = pack_bf16(__uint_as_float(ra[g * 8 + 2 * i]) * f, __uint_as_float(ra[g * 8 + 2 * i + 1]) * f);
            wb[i] = pack_bf16(__uint_as_float(rb[g * 8 + 2 * i]) * f, __uint_as_float(rb[g * 8 + 2 * i + 1]) * f);
          }
          *reinterpret_cast<uint4*>(orow + ((g ^ (row & 7)) << 4)) = make_uint4(wa[0], wa[1], wa[2], wa[3]);
          *reinterpret_cast<uint4*>(orow + (((4 + g) ^ (row & 7)) << 4)) = make_uint4(wb[0], wb[1], wb[2], wb[3]);
        }
      }
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(smem_u32(&S.dkv_empty));
        mbar_arrive(smem_u32(&S.dkv_staged));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

}  // namespace
}  // namespace yat

// ------------------------------------------------------------------ host side
YAT_DEFINE_TRACE_API(yat_xbwd)

extern "C" int64_t yat_xattn_bwd_workspace_bytes(int B, int H, int Nq) { return (int64_t)B * H * Nq * (int64_t)sizeof(float); }

extern "C" int yat_xattn_bwd(const void* q, const void* k, const void* v, const void* bias, const void* o, const void* lse,
                             const void* d_o, int B, int H, int Nq, int L, int d, int64_t ldq, int64_t ldkv, float scale,
                             void* dq, void* dk, void* dv, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(q && k && v && o && lse && d_o && dq && workspace, YAT_ERR_BAD_ARG, "yat_xattn_bwd: null pointer argument");
  YAT_REQUIRE((dk == nullptr) == (dv == nullptr), YAT_ERR_BAD_ARG, "yat_xattn_bwd: dk and dv go together");
  YAT_REQUIRE(B > 0 && H > 0 && Nq > 0 && L > 0, YAT_ERR_BAD_ARG, "yat_xattn_bwd: non-positive dimension");
  YAT_REQUIRE(d >= 8 && d <= 128 && d % 8 == 0, YAT_ERR_UNSUPPORTED, "yat_xattn_bwd: head_dim %d not supported (8..128, multiple of 8)", d);
  YAT_REQUIRE(L <= kMaxKeys, YAT_ERR_UNSUPPORTED, "yat_xattn_bwd: %d keys exceed the %d supported", L, kMaxKeys);
  YAT_REQUIRE((int64_t)H * d <= 12288, YAT_ERR_UNSUPPORTED, "yat_xattn_bwd: H * d = %lld exceeds 12288", (long long)H * d);
  YAT_REQUIRE(ldq >= (int64_t)H * d && ldq % 8 == 0 && ldkv >= (int64_t)H * d && ldkv % 8 == 0, YAT_ERR_BAD_ARG,
              "yat_xattn_bwd: row pitches %lld / %lld must be >= H*d and multiples of 8", (long long)ldq, (long long)ldkv);
  YAT_REQUIRE((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)o | (uintptr_t)d_o | (uintptr_t)dq | (uintptr_t)dk |
                (uintptr_t)dv) & 15u) == 0, YAT_ERR_BAD_ARG, "yat_xattn_bwd: tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rows = (int64_t)B * Nq;
  float* D = (float*)workspace;
  xattn_prep_kernel<<<(unsigned)((rows + kPrepRows - 1) / kPrepRows), 256, (size_t)kPrepRows * H * d / 8 * sizeof(float), st>>>(
      (const __nv_bfloat16*)o, (const __nv_bfloat16*)d_o, D, rows, H, d, Nq);
  int rc = check_launch("xattn_prep");
  if (rc != YAT_OK) return rc;
  const uint64_t C = (uint64_t)H * d;
  CUtensorMap tq, tdo, tk, tv, tdq;
  if ((rc = make_tmap_heads_bf16(&tq, q, d, H, Nq, B, (uint64_t)ldq, kQT)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tdo, d_o, d, H, Nq, B, C, kQT)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tk, k, d, H, L, B, (uint64_t)ldkv, kKR)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tv, v, d, H, L, B, (uint64_t)ldkv, kKR)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tdq, dq, d, H, Nq, B, C, kQT)) != YAT_OK) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(xattn_dq_sm100_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDqSmemBytes);
    YAT_REQUIRE(e == cudaSuccess, YAT_ERR_CUDA, "yat_xattn_bwd: cannot reserve %d B of shared memory (%s)", kDqSmemBytes,
                cudaGetErrorString(e));
    attr_set = true;
  }
  const int items = B * H * ((Nq + kQT - 1) / kQT);
  xattn_dq_sm100_kernel<<<items < kNumSMs ? items : kNumSMs, kThreads, kDqSmemBytes, st>>>(
      tq, tdo, tk, tv, tdq, (const float*)bias, (const float*)lse, D, B, H, Nq, L, d, scale);
  rc = check_launch("xattn_dq_sm100");
  if (rc != YAT_OK) return rc;
  if (dk == nullptr) return YAT_OK;
  CUtensorMap tq2, tdo2, tk2, tv2, tdk, tdv;      // the dkv kernel reads 128-row boxes of everything
  if ((rc = make_tmap_heads_bf16(&tq2, q, d, H, Nq, B, (uint64_t)ldq, kQT)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tdo2, d_o, d, H, Nq, B, C, kQT)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tk2, k, d, H, L, B, (uint64_t)ldkv, kKB)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tv2, v, d, H, L, B, (uint64_t)ldkv, kKB)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tdk, dk, d, H, L, B, C, kKB)) != YAT_OK) return rc;
  if ((rc = make_tmap_heads_bf16(&tdv, dv, d, H, L, B, C, kKB)) != YAT_OK) return rc;
  static bool attr2_set = false;
  if (!attr2_set) {
    cudaError_t e = cudaFuncSetAttribute(xattn_dkv_sm100_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDkvSmemBytes);
    YAT_REQUIRE(e == cudaSuccess, YAT_ERR_CUDA, "yat_xattn_bwd: cannot reserve %d B of shared memory (%s)", kDkvSmemBytes,
                cudaGetErrorString(e));
    attr2_set = true;
  }
  const int kitems = B * H * ((L + kKB - 1) / kKB);
  xattn_dkv_sm100_kernel<<<kitems < kNumSMs ? kitems : kNumSMs, kThreads, kDkvSmemBytes, st>>>(
      tq2, tdo2, tk2, tv2, tdk, tdv, (const float*)bias, (const float*)lse, D, B, H, Nq, L, d, scale);
  return check_launch("xattn_dkv_sm100");
}
