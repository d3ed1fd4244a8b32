// SANA ReLU linear attention on the 5th-gen tensor cores (tcgen05 + TMEM + TMA), forward.
//
//   O_n = relu(Q_n) (relu(K)^T V) / (relu(Q_n) . sum_m relu(K_m) + 1e-15)          per (batch, head)
//
// replaces diffusers' SanaLinearAttnProcessor2_0 core (transposes, .float() copies, F.pad and two
// fp32 SIMT batched matmuls; reached from utils/patch_sana_attention_layers.py:94).
//
// Layout: q,k,v,o are [B, N, H*32] bf16 row-major, head h = columns [32h, 32h+32).  A work item is
// (batch b, group of 4 heads) = a 128-column stripe; it is streamed in 128-token tiles of
// [2 halves][128 tokens][64 cols] (two SWIZZLE_128B TMA boxes, 32 KB) through one shared-memory ring:
//   K0 V0 K1 V1 ... | Q0 Q1 ...         (3 * ceil(N/128) tiles per item; O tiles leave from the Q slot)
// The grid is persistent (one CTA per SM, loads run ahead across items).
//
// Roles (608 threads):
//   warp 0       TMA producer
//   warp 1       tcgen05 issuer (one thread)
//   warps 2-9    transform/epilogue group 0   } phase 2: the two groups own alternate tiles, so the ReLU pass of
//   warps 10-17  transform/epilogue group 1   } tile t+1 overlaps the MMA + epilogue of tile t (2 heads per thread)
//   warp 18      TMA-store warp: O tile out, then the slot goes back to the producer
// In phase 1 all sixteen warps ReLU every K tile together, and Ks = relu(K)^T 1 is a 16-column MMA against a ones
// operand (round 2: the CUDA-core column sums + their CTA barrier made the K transform the longest stage of the ring).
//
//   phase 1  transform: relu(K) in place     MMA1: D1[128x128] += relu(K)^T V,  Ks[128x16] += relu(K)^T ones
//            (A and B both MN-major straight from the TMA tiles; the four 32x32 diagonal blocks of D1
//             are the per-head K^T V, the off-diagonal blocks are discarded — the kernel is HBM-bound)
//   build    (group 0) D1 diagonal -> registers -> bf16 hi + bf16 lo split -> smem B operand
//            [32 x 48] per head: columns 0..31 = KV_h, column 32 = Ks_h, rest 0.  hi + lo keeps the
//            second contraction at ~fp32 accuracy (the reference does both matmuls in fp32).
//   phase 2  transform: relu(Q) in place          MMA2 per head: D2[128 x 48] = relu(Q_h) [KV_h | Ks_h]
//            (K-major A, MN-major B): columns 0..31 are the numerator, column 32 the normaliser
//            epilogue: TMEM -> registers, divide, bf16, swizzled smem (the Q slot) -> store warp.
// Rows beyond N and heads beyond H are zero-filled by TMA on load and clipped on store.
#include "common.cuh"
#include "sm100.cuh"

namespace yat {
namespace {
using namespace sm100;

constexpr int kD = 32;
constexpr int kHeadsPerItem = 4;
constexpr int kTile = 128;                       // tokens per tile
constexpr int kHalfBytes = kTile * 128;          // [128 tokens][64 cols] bf16
constexpr int kSlotBytes = 2 * kHalfBytes;       // 32 KB
constexpr int kStages = 6;
constexpr int kGroups = 2, kGroupWarps = 8;       // two groups of 8 warps: 2 warps per TMEM lane quadrant, 2 heads each
constexpr int kEpiWarps = kGroups * kGroupWarps;
constexpr int kStoreWarp = 2 + kEpiWarps;
constexpr int kThreads = 32 * (kStoreWarp + 1);
constexpr int kNB = 48;                          // N of the phase-2 MMA: 32 KV columns + Ks + padding to 16
constexpr int kD2Cols = kHeadsPerItem * kNB;     // 192 TMEM columns per D2 buffer
constexpr int kTmemCols = 512;                   // D1 [0,128) + D2[group] [128 + 192 g, +192)
constexpr int kBmatBytes = kD * kNB * 2;         // one [32 x 48] bf16 B operand
constexpr int kVecBytes = kTile * 16 * 2;        // [128 tokens x 16] MN-major ones operand: relu(K)^T 1 = Ks on the tensor core
constexpr uint32_t kColD2 = 128;                 // D2[group] at kColD2 + group * kD2Cols
constexpr uint32_t kColKs = kColD2;              // phase 1: the Ks accumulator borrows the first 16 columns of D2[0]

struct FwdSmem {
  uint8_t slots[kStages][kSlotBytes];                       // 1024-aligned
  uint8_t bmat[kHeadsPerItem][2][kBmatBytes];               // [head][hi|lo], MN-major no-swizzle
  uint8_t ones[kVecBytes];
  uint64_t full[kStages], ready[kStages], ready_k[kStages], empty[kStages], o_ready[kStages];
  uint64_t d1_full, b_ready, d2_full[kGroups], d2_empty[kGroups];
  uint32_t tmem_base;
};
constexpr int kFwdSmemBytes = (int)sizeof(FwdSmem) + 1024;
static_assert(kFwdSmemBytes <= 232448, "forward shared memory exceeds the 227 KB per-CTA limit");

constexpr uint32_t kIdesc1 = make_idesc_bf16(128, 128, 1, 1);  // relu(K)^T V : A, B MN-major
constexpr uint32_t kIdesc2 = make_idesc_bf16(128, kNB, 0, 1);  // relu(Q) [KV|Ks] : A K-major, B MN-major
constexpr uint32_t kIdescVec = make_idesc_bf16(128, 16, 1, 1); // relu(K)^T ones over tokens

__device__ __forceinline__ uint4 relu8(uint4 v) {
  const __nv_bfloat162 z = __float2bfloat162_rn(0.f);
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __hmax2_nan(h[i], z);
  return v;
}

// kState: phase 1 has already been done by the projection kernel (qkv_linattn_sm100.cu): KV_h / Ks_h arrive as fp32
// partials state[b][tile][head][33][32] (n_part tiles per sample, summed here in fixed order) and only q is streamed.
template <bool kState>
__global__ void __launch_bounds__(kThreads, 1)
linattn_fwd_sm100_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                         const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_o,
                         int B, int N, int groups, const float* __restrict__ state, int n_part, int H) {
  extern __shared__ uint8_t smem_raw[];
  FwdSmem& S = *reinterpret_cast<FwdSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = (N + kTile - 1) / kTile;
  const int n_items = B * groups;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(smem_u32(&S.full[s]), 1);
      mbar_init(smem_u32(&S.ready[s]), kGroupWarps);      // phase 2: the owning group's warps
      mbar_init(smem_u32(&S.ready_k[s]), kEpiWarps);      // phase 1: every epilogue warp works on every K tile
      mbar_init(smem_u32(&S.empty[s]), 1);
      mbar_init(smem_u32(&S.o_ready[s]), kGroupWarps);
    }
    mbar_init(smem_u32(&S.d1_full), 1);
    mbar_init(smem_u32(&S.b_ready), 4);                   // the four warps (one per head) that build the operands
    for (int i = 0; i < kGroups; ++i) {
      mbar_init(smem_u32(&S.d2_full[i]), 1);
      mbar_init(smem_u32(&S.d2_empty[i]), kGroupWarps);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tm_q); tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v); tma_prefetch_desc(&tm_o);
  }
  // the padding columns of the B operands stay zero for the whole kernel; ones operand: column 0 = 1
  for (int i = threadIdx.x; i < (int)(sizeof(S.bmat) + sizeof(S.ones)) / 16; i += kThreads)
    reinterpret_cast<uint4*>(&S.bmat[0][0][0])[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  if (threadIdx.x < kTile)   // element [token][n] at (n/8)*2048 + (token/8)*128 + (token%8)*16 + (n%8)*2
    *reinterpret_cast<uint32_t*>(&S.ones[(threadIdx.x >> 3) * 128 + (threadIdx.x & 7) * 16]) = 0x00003f80u;   // bf16 1.0
  fence_proxy_async();
  if (warp == 1) {
    tmem_alloc(smem_u32(&S.tmem_base), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // one CTA per SM allocating all 512 columns: the base is column 0 (immediates in the issuer's MMA loop, see the backward)
  if (S.tmem_base != 0) __trap();
  constexpr uint32_t tmem = 0;

  if (warp == 0) {
    // ======================= TMA producer =======================
    if (lane == 0) {
      uint32_t it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int b = item / groups, c0 = (item % groups) * (kHeadsPerItem * kD);
        for (int j = kState ? 2 * n_tiles : 0; j < 3 * n_tiles; ++j, ++it) {
          const CUtensorMap* tm;
          int tile;
          if (j < 2 * n_tiles) { tm = (j & 1) ? &tm_v : &tm_k; tile = j >> 1; }
          else                 { tm = &tm_q; tile = j - 2 * n_tiles; }
          const uint32_t s = it % kStages, ph = (it / kStages) & 1;
          mbar_wait(smem_u32(&S.empty[s]), ph ^ 1);
          const uint32_t full = smem_u32(&S.full[s]);
          mbar_arrive_expect_tx(full, kSlotBytes);
          const uint32_t dst = smem_u32(&S.slots[s][0]);
          tma_load_3d(dst, tm, full, c0, tile * kTile, b);
          tma_load_3d(dst + kHalfBytes, tm, full, c0 + 64, tile * kTile, b);
        }
      }
    }
  } else if (warp == 1) {
    // ======================= tcgen05 issuer =======================
    if (lane == 0) {
      uint32_t it = 0, item_iter = 0;
      uint32_t ready_ph = 0, ready_k_ph = 0;   // per-slot parities of `ready` (Q uses) / `ready_k` (K uses)
      uint32_t d2_use[kGroups] = {0, 0};
      const uint64_t ones0 = make_smem_desc(smem_u32(&S.ones[0]), 128, 2048, kLayoutNone);
      const uint64_t bmat0 = make_smem_desc(smem_u32(&S.bmat[0][0][0]), 128, 512, kLayoutNone);
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++item_iter) {
        // ---- phase 1: D1 += relu(K)^T V over all token tiles
        for (int t = 0; t < (kState ? 0 : n_tiles); ++t, it += 2) {
          const uint32_t sk = it % kStages;
          const uint32_t sv = (it + 1) % kStages, pv = ((it + 1) / kStages) & 1;
          TRACE(1, 50);
          mbar_wait(smem_u32(&S.ready_k[sk]), (ready_k_ph >> sk) & 1);
          ready_k_ph ^= 1u << sk;
          TRACE(1, 51);
          mbar_wait(smem_u32(&S.full[sv]), pv);  // V is consumed as loaded
          // the Ks accumulator borrows the first columns of D2[0]: the previous item's last group-0 tile must be drained
          if (t == 0) mbar_wait(smem_u32(&S.d2_empty[0]), (d2_use[0] & 1) ^ 1);
          tc_fence_after();
          TRACE(1, 52);
          // MN-major SW128: LBO = stride between the two 64-column halves, SBO = 8-token group (1 KB); a descriptor is
          // its base + (byte offset >> 4)
          const uint64_t ka0 = make_smem_desc(smem_u32(&S.slots[sk][0]), kHalfBytes, 1024, kLayoutSw128);
          const uint64_t va0 = make_smem_desc(smem_u32(&S.slots[sv][0]), kHalfBytes, 1024, kLayoutSw128);
#pragma unroll
          for (int ks = 0; ks < kTile / 16; ++ks)
            umma_bf16(tmem, ka0 + ((ks * 2048) >> 4), va0 + ((ks * 2048) >> 4), kIdesc1, (t | ks) != 0);
#pragma unroll
          for (int ks = 0; ks < kTile / 16; ++ks)    // Ks = relu(K)^T 1 (column 0 of a 16-column accumulator)
            umma_bf16(tmem + kColKs, ka0 + ((ks * 2048) >> 4), ones0 + ((ks * 256) >> 4), kIdescVec, (t | ks) != 0);
          tc_commit(smem_u32(&S.empty[sk]));
          tc_commit(smem_u32(&S.empty[sv]));
          TRACE(1, 53);
        }
        if (!kState) tc_commit(smem_u32(&S.d1_full));
        mbar_wait(smem_u32(&S.b_ready), item_iter & 1);
        tc_fence_after();
        // ---- phase 2: D2[group] = relu(Q_h) ([KV_h|Ks_h] hi + lo), per head
        for (int t = 0; t < n_tiles; ++t, ++it) {
          const uint32_t sq = it % kStages;
          const uint32_t g = t & 1, pb = d2_use[g] & 1;
          ++d2_use[g];
          TRACE(1, 30);
          mbar_wait(smem_u32(&S.ready[sq]), (ready_ph >> sq) & 1);
          ready_ph ^= 1u << sq;
          TRACE(1, 31);
          mbar_wait(smem_u32(&S.d2_empty[g]), pb ^ 1);
          tc_fence_after();
          TRACE(1, 32);
          // A: K-major SW128, 8-row groups 1 KB apart; head h = 64 B inside the 128 B row of half h >> 1; K step = 32 B
          const uint64_t qa0 = make_smem_desc(smem_u32(&S.slots[sq][0]), 16, 1024, kLayoutSw128);
#pragma unroll
          for (int h = 0; h < kHeadsPerItem; ++h) {
            const uint32_t d2 = tmem + kColD2 + g * kD2Cols + h * kNB;
            const uint32_t aoff = (h >> 1) * kHalfBytes + (h & 1) * 64;
#pragma unroll
            for (int hl = 0; hl < 2; ++hl) {
              const uint32_t boff = (h * 2 + hl) * kBmatBytes;
#pragma unroll
              for (int ks = 0; ks < 2; ++ks)
                // B: MN-major no-swizzle, core matrix = 8 k-rows x 16 B; k-group stride (LBO) 128 B,
                //    mn-group stride (SBO) 512 B; K step of 16 = 2 k-groups = 256 B
                umma_bf16(d2, qa0 + ((aoff + ks * 32) >> 4), bmat0 + ((boff + ks * 256) >> 4), kIdesc2, (hl | ks) != 0);
            }
          }
          tc_commit(smem_u32(&S.d2_full[g]));
          TRACE(1, 33);
        }
        if (kState) tc_commit(smem_u32(&S.d1_full));   // every MMA of this item has read bmat: it may be rebuilt
      }
    }
  } else if (warp == kStoreWarp) {
    // ======================= TMA-store warp =======================
    if (lane == 0) {
      uint32_t it = 0, o_ph = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int b = item / groups, c0 = (item % groups) * (kHeadsPerItem * kD);
        if (!kState) it += 2 * n_tiles;
        for (int t = 0; t < n_tiles; ++t, ++it) {
          const uint32_t sq = it % kStages;
          TRACE(2, 40);
          mbar_wait(smem_u32(&S.o_ready[sq]), (o_ph >> sq) & 1);
          o_ph ^= 1u << sq;
          TRACE(2, 41);
          const uint32_t src = smem_u32(&S.slots[sq][0]);
          tma_store_3d(&tm_o, src, c0, t * kTile, b);
          tma_store_3d(&tm_o, src + kHalfBytes, c0 + 64, t * kTile, b);
          tma_store_commit();
          tma_store_wait_read<0>();                       // smem has been read: the slot can be refilled
          mbar_arrive(smem_u32(&S.empty[sq]));
          TRACE(2, 42);
        }
      }
      tma_store_wait_all<0>();
    }
  } else {
    // ======================= transform + epilogue groups =======================
    const int te = threadIdx.x - 64;          // 0..511
    const int we = warp - 2;                  // 0..15
    const int grp = we >> 3;                  // phase 2: tiles with (t & 1) == grp belong to this group
    const int tg = te - grp * (kGroupWarps * 32);  // 0..255 inside the group
    const int cc = te & 15;                   // logical 16-byte chunk column of the 256-byte item row (te and tg agree mod 16)
    const int half = cc >> 3, chunk = cc & 7;
    const int quad = warp & 3;                // TMEM lane quadrant this warp may access
    const int hp = (we >> 2) & 1;             // phase-2 epilogue: heads 2 hp, 2 hp + 1 of this thread's row
    const int row = quad * 32 + lane;         // token row in the tile (TMEM lane)
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    uint32_t it = 0, item_iter = 0, d2_use = 0;

    // relu in place: rows r0, r0 + step, ... (n rows per thread) of a [2][128][64] tile, then one arrival per warp
    auto relu_rows = [&](uint32_t slot, int r0, int step, int n, uint64_t* bar) {
      uint8_t* base = &S.slots[slot][half * kHalfBytes];
#pragma unroll 8
      for (int i = 0; i < n; ++i) {
        const int r = r0 + step * i;
        uint4* p = reinterpret_cast<uint4*>(base + r * 128 + ((chunk ^ (r & 7)) << 4));
        *p = relu8(*p);
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(bar));
    };

    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++item_iter) {
      // ---- phase 1: relu(K) in place by ALL sixteen warps (4 rows per thread); the column sums Ks come off the tensor
      // core (relu(K)^T ones), so there is no CUDA-core reduction and no CTA-wide barrier here.  Every `full` phase is
      // observed in order (a parity wait is only sound if the waiter sees every phase of the barrier).  V needs no
      // transform: the issuer waits on its `full`.
      for (int t = 0; t < (kState ? 0 : n_tiles); ++t, it += 2) {
        const uint32_t sk = it % kStages, pk = (it / kStages) & 1;
        if (tg == 0) TRACE(grp ? 3 : 0, 20);
        mbar_wait(smem_u32(&S.full[sk]), pk);
        if (tg == 0) TRACE(grp ? 3 : 0, 21);
        relu_rows(sk, te >> 4, 32, 4, &S.ready_k[sk]);
        if (tg == 0) TRACE(grp ? 3 : 0, 22);
        mbar_wait(smem_u32(&S.full[(it + 1) % kStages]), ((it + 1) / kStages) & 1);
        if (tg == 0) TRACE(grp ? 3 : 0, 23);
      }
      if (we < 4) {
        // ---- build the B operands of phase 2: head `quad`, k row `lane`
        float ksum = 0.f;
        uint32_t r[32];
        if (kState) {
          // the previous item's phase-2 MMAs read bmat: the issuer commits d1_full after the last of them (kState only)
          if (item_iter > 0) mbar_wait(smem_u32(&S.d1_full), (item_iter - 1) & 1);
          const int head = (item % groups) * kHeadsPerItem + quad;
          float accv[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) accv[j] = 0.f;
          if (head < H) {
            const float* sp = state + (((int64_t)(item / groups) * n_part) * H + head) * (33 * 32);
            for (int p = 0; p < n_part; ++p, sp += (int64_t)H * (33 * 32)) {
              const float4* row4 = reinterpret_cast<const float4*>(sp + lane * 32);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 v = __ldg(row4 + j);
                accv[4 * j] += v.x; accv[4 * j + 1] += v.y; accv[4 * j + 2] += v.z; accv[4 * j + 3] += v.w;
              }
              ksum += __ldg(sp + 32 * 32 + lane);
            }
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(accv[j]);
        } else {
          mbar_wait(smem_u32(&S.d1_full), item_iter & 1);
          tc_fence_after();
          uint32_t k1;
          tmem_ld32(tmem + lane_base + quad * kD, r);  // KV_head[dk = lane][dv = 0..31]
          tmem_ld1(tmem + lane_base + kColKs, k1);     // Ks[column quad * 32 + lane]
          tmem_ld_wait();
          ksum = __uint_as_float(k1);
        }
        uint8_t* bh = &S.bmat[quad][0][0];
        uint8_t* bl = &S.bmat[quad][1][0];
        const int off = (lane >> 3) * 128 + (lane & 7) * 16;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t hi[4], lo[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float a0 = __uint_as_float(r[g * 8 + 2 * j]), a1 = __uint_as_float(r[g * 8 + 2 * j + 1]);
            hi[j] = pack_bf16(a0, a1);
            lo[j] = pack_bf16(a0 - bf16_lo(hi[j]), a1 - bf16_hi(hi[j]));
          }
          *reinterpret_cast<uint4*>(bh + g * 512 + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          *reinterpret_cast<uint4*>(bl + g * 512 + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        {  // column 32 = Ks (mn-group 4, first element); the other 7 elements of the chunk stay zero
          const uint32_t khi = pack_bf16(ksum, 0.f);
          const uint32_t klo = pack_bf16(ksum - bf16_lo(khi), 0.f);
          *reinterpret_cast<uint4*>(bh + 4 * 512 + off) = make_uint4(khi, 0, 0, 0);
          *reinterpret_cast<uint4*>(bl + 4 * 512 + off) = make_uint4(klo, 0, 0, 0);
        }
        fence_proxy_async();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.b_ready));
      }

      // ---- phase 2
      for (int t = 0; t < n_tiles; ++t, ++it) {
        const uint32_t sq = it % kStages, pq = (it / kStages) & 1;
        if (tg == 0) TRACE(grp ? 3 : 0, 9);
        mbar_wait(smem_u32(&S.full[sq]), pq);
        if ((t & 1) != grp) continue;
        if (tg == 0) TRACE(grp ? 3 : 0, 10);
        relu_rows(sq, tg >> 4, 16, 8, &S.ready[sq]);
        if (tg == 0) TRACE(grp ? 3 : 0, 11);

        mbar_wait(smem_u32(&S.d2_full[grp]), d2_use & 1);
        ++d2_use;
        tc_fence_after();
        if (tg == 0) TRACE(grp ? 3 : 0, 12);
        const uint32_t d2 = tmem + lane_base + kColD2 + grp * kD2Cols;
#pragma unroll
        for (int h = 2 * hp; h < 2 * hp + 2; ++h) {
          uint32_t acc[32], dn;
          tmem_ld32(d2 + h * kNB, acc);
          tmem_ld1(d2 + h * kNB + kD, dn);
          tmem_ld_wait();
          const float rden = __frcp_rn(__uint_as_float(dn) + 1e-15f);
          uint8_t* orow = &S.slots[sq][(h >> 1) * kHalfBytes + row * 128];
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint32_t w[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
              w[j] = pack_bf16(__uint_as_float(acc[g * 8 + 2 * j]) * rden, __uint_as_float(acc[g * 8 + 2 * j + 1]) * rden);
            const int c = (h & 1) * 4 + g;
            *reinterpret_cast<uint4*>(orow + ((c ^ (row & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
          }
        }
        tc_fence_before();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(smem_u32(&S.d2_empty[grp]));
          mbar_arrive(smem_u32(&S.o_ready[sq]));
        }
        if (tg == 0) TRACE(grp ? 3 : 0, 13);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

}  // namespace

// ------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  }
  return fn;
}

int make_tmap_3d_bf16(CUtensorMap* out, const void* base, uint64_t C, uint64_t N, uint64_t B, uint32_t box_c,
                      uint32_t box_n, uint64_t ld) {
  PFN_encodeTiled enc = get_encode_fn();
  YAT_REQUIRE(enc != nullptr, YAT_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  if (ld == 0) ld = C;
  const cuuint64_t dims[3] = {C, N, B};
  const cuuint64_t strides[2] = {ld * 2, N * ld * 2};
  const cuuint32_t box[3] = {box_c, box_n, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    // first driver-API call on this thread (autograd's backward worker): bind the primary context and retry
    cudaFree(nullptr);
    r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  YAT_REQUIRE(r == CUDA_SUCCESS, YAT_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return YAT_OK;
}

int make_tmap_heads_bf16(CUtensorMap* out, const void* base, uint64_t d, uint64_t H, uint64_t rows, uint64_t B, uint64_t ld,
                         uint32_t box_rows) {
  PFN_encodeTiled enc = get_encode_fn();
  YAT_REQUIRE(enc != nullptr, YAT_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  const cuuint64_t dims[4] = {d, H, rows, B};
  const cuuint64_t strides[3] = {d * 2, ld * 2, rows * ld * 2};
  const cuuint32_t box[4] = {64, 1, box_rows, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = CUDA_ERROR_UNKNOWN;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT) break;
    cudaFree(nullptr);  // first driver-API call on this thread (autograd worker): bind the primary context, retry
  }
  YAT_REQUIRE(r == CUDA_SUCCESS, YAT_ERR_CUDA, "cuTensorMapEncodeTiled (heads) failed with CUresult %d", (int)r);
  return YAT_OK;
}

// Plain (unswizzled) 4-D bf16 tensor map {d0, d1, d2, d3} (d0 contiguous) with box {b0, b1, b2, 1}; out-of-bounds
// elements of a box (negative or past-the-end coordinates) read as zero.
int make_tmap_4d_bf16_plain(CUtensorMap* out, const void* base, const uint64_t dims[4], const uint32_t box[4]) {
  PFN_encodeTiled enc = get_encode_fn();
  YAT_REQUIRE(enc != nullptr, YAT_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  const cuuint64_t d[4] = {dims[0], dims[1], dims[2], dims[3]};
  const cuuint64_t strides[3] = {dims[0] * 2, dims[0] * dims[1] * 2, dims[0] * dims[1] * dims[2] * 2};
  const cuuint32_t bx[4] = {box[0], box[1], box[2], box[3]};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = CUDA_ERROR_UNKNOWN;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), d, strides, bx, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT) break;
    cudaFree(nullptr);  // first driver-API call on this thread (autograd worker): bind the primary context, retry
  }
  YAT_REQUIRE(r == CUDA_SUCCESS, YAT_ERR_CUDA, "cuTensorMapEncodeTiled (4d) failed with CUresult %d", (int)r);
  return YAT_OK;
}

}  // namespace yat

YAT_DEFINE_TRACE_API(yat_fwd)

extern "C" int yat_linattn_fwd(const void* q, const void* k, const void* v, int B, int N, int H, int D, void* o,
                               void* stream) {
  return yat_linattn_fwd_ld(q, k, v, (int64_t)H * D, B, N, H, D, o, stream);
}

static int linattn_fwd_launch(const void* q, const void* k, const void* v, int64_t ld, int B, int N, int H, int D, void* o,
                              const float* state, int n_part, void* stream) {
  using namespace yat;
  YAT_REQUIRE(q && k && v && o, YAT_ERR_BAD_ARG, "yat_linattn_fwd: null pointer argument");
  YAT_REQUIRE(ld >= (int64_t)H * D && ld % 8 == 0, YAT_ERR_BAD_ARG, "yat_linattn_fwd: row pitch %lld must be >= H*D and a multiple of 8",
              (long long)ld);
  YAT_REQUIRE(B > 0 && N > 0 && H > 0, YAT_ERR_BAD_ARG, "yat_linattn_fwd: non-positive dimension");
  YAT_REQUIRE(D == kD, YAT_ERR_UNSUPPORTED, "yat_linattn_fwd: head_dim %d not supported (SANA uses 32)", D);
  YAT_REQUIRE((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)o | (uintptr_t)state) & 15u) == 0, YAT_ERR_BAD_ARG,
              "yat_linattn_fwd: tensors must be 16-byte aligned");
  CUtensorMap tq, tk, tv, to;
  const uint64_t C = (uint64_t)H * D;
  int rc;
  if ((rc = make_tmap_3d_bf16(&tq, q, C, N, B, 64, kTile, (uint64_t)ld)) != YAT_OK) return rc;
  if ((rc = make_tmap_3d_bf16(&tk, k, C, N, B, 64, kTile, (uint64_t)ld)) != YAT_OK) return rc;
  if ((rc = make_tmap_3d_bf16(&tv, v, C, N, B, 64, kTile, (uint64_t)ld)) != YAT_OK) return rc;
  if ((rc = make_tmap_3d_bf16(&to, o, C, N, B, 64, kTile)) != YAT_OK) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(linattn_fwd_sm100_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmemBytes);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(linattn_fwd_sm100_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmemBytes);
    YAT_REQUIRE(e == cudaSuccess, YAT_ERR_CUDA, "yat_linattn_fwd: cannot reserve %d B of shared memory (%s)",
                kFwdSmemBytes, cudaGetErrorString(e));
    attr_set = true;
  }
  const int groups = (H + kHeadsPerItem - 1) / kHeadsPerItem;
  const int items = B * groups;
  const int grid = items < kNumSMs ? items : kNumSMs;
  if (state)
    linattn_fwd_sm100_kernel<true><<<grid, kThreads, kFwdSmemBytes, (cudaStream_t)stream>>>(tq, tk, tv, to, B, N, groups, state, n_part, H);
  else
    linattn_fwd_sm100_kernel<false><<<grid, kThreads, kFwdSmemBytes, (cudaStream_t)stream>>>(tq, tk, tv, to, B, N, groups, nullptr, 0, H);
  return check_launch("linattn_fwd_sm100");
}

extern "C" int yat_linattn_fwd_ld(const void* q, const void* k, const void* v, int64_t ld, int B, int N, int H, int D,
                                  void* o, void* stream) {
  return linattn_fwd_launch(q, k, v, ld, B, N, H, D, o, nullptr, 0, stream);
}

// Phase 2 only: q [B, N, H*32] (row pitch ld) and the state partials written by yat_qkv_linattn_fwd
// (state[b][tile][head][33][32] fp32, 2 * ceil(N / 256) tiles per sample) -> o [B, N, H*32].
extern "C" int yat_linattn_fwd_state(const void* q, int64_t ld, const void* state, int B, int N, int H, int D, void* o,
                                     void* stream) {
  YAT_REQUIRE(state != nullptr, YAT_ERR_BAD_ARG, "yat_linattn_fwd_state: null state");
  return linattn_fwd_launch(q, q, q, ld, B, N, H, D, o, (const float*)state, 2 * ((N + 255) / 256), stream);
}
