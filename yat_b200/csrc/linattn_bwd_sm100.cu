// SANA ReLU linear attention, backward, on the 5th-gen tensor cores (tcgen05 + TMEM + TMA).
//
// Per (batch, head), Q = relu(q), K = relu(k), D = 32 (derivation: SURVEY.md §8 A10, oracle/linear_attention.py):
//   KV = K^T V   Ks = sum_n K_n   den_n = Q_n.Ks + 1e-15   num_n = Q_n KV   r_n = 1/den_n
//   dDen_n = -(dO_n . num_n) r_n^2
//   dQ_n  = (r_n (dO_n KV^T) + dDen_n Ks) * [q_n > 0]
//   dKV   = sum_n (r_n Q_n)^T dO_n          dKs = sum_n Q_n dDen_n
//   dK_n  = (V_n dKV^T + dKs) * [k_n > 0]    dV_n = K_n dKV
// Every reduction over tokens or over a head's 32 columns runs on the tensor core (the column sums Ks and dKs as
// 16-column MMAs against a ones / per-row-weight operand, den as a 33rd column of the KV operand): the CUDA cores only
// do elementwise work on their own rows, so pass B has no cross-thread shared-memory traffic and no CTA-wide barrier.
// Nothing but q, k, v, dO is read (O is recomputed: it is what gradient checkpointing would do anyway),
// so the algorithmic traffic is 7 tensors; k and v are streamed twice (pass A and pass C), the second time
// mostly out of L2.
//
// Work item = (batch, 4 heads) = a 128-column stripe, streamed in 128-token tiles through a ring of 32 KB
// shared-memory slots ([2 halves][128 tokens][64 cols], SWIZZLE_128B), exactly like the forward kernel:
//   pass A   K V  pairs -> D_KV [128x128] += relu(K)^T V (diagonal 32x32 blocks = per-head KV), Ks = relu(K)^T 1
//   pass B   Q dO pairs -> [num | den] = Q [KV | Ks] -> r, dDen; Q~ = bf16(r Q) in place;
//            dQ~ = dO KV^T (raw dO, exact) ; D_dKV += Q~^T dO ; D_dKs += Q~^T (dDen / r) ; dQ leaves from the dO slot
//   pass C   K V  pairs -> dV = relu(K) dKV, dK = V dKV^T + dKs ; both leave from their own slots
// The 32x32 KV / dKV operands are split into bf16 hi + lo and the per-row scales of dQ are applied in the fp32
// epilogue, so the contractions whose operand is an ACCUMULATED quantity keep ~fp32 accuracy like the reference's
// `.float()` matmuls.  Q~ = r Q is rounded to bf16 once: its rounding errors are independent
// per token and average out over the N-token sum of dKV / dKs (measured: dq, dk, dv at 1.66e-3 of the fp64
// oracle at the SANA-1.6B size with or without a lo half, i.e. the bf16 rounding of the outputs themselves), and
// the 32 KB the lo half occupied is the sixth ring slot (165 -> 154 us).
//
// Roles (608 threads): warp 0 TMA producer, warp 1 tcgen05 issuer, warps 2..17 transform/epilogue (16 warps: the
// CUDA-core work between the MMAs is what bounds the kernel, and 8 warps left the schedulers idle 70 % of the
// time), warp 18 TMA-store + slot release.  Passes B and C are software-pipelined by one tile: the epilogue warps
// transform tile t+1 (relu) before they finish tile t, the issuer orders the tensor pipe dQ~(t), dKV(t) + dKs(t),
// num(t+1), and pass C double-buffers its dV / dK accumulators in TMEM - the MMA + mbarrier round trips that used
// to sit between E1 -> E2 -> E3 of one tile now overlap the neighbouring tile's CUDA-core work.  Parity waits follow one rule: every waiter observes every phase of
// the barriers it waits on, in order.
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "sm100.cuh"

namespace yat {
namespace {
using namespace sm100;

constexpr int kD = 32;
constexpr int kHeadsPerItem = 4;
constexpr int kTile = 128;
constexpr int kHalfBytes = kTile * 128;
constexpr int kSlotBytes = 2 * kHalfBytes;
constexpr int kStages = 6;
constexpr int kEpiWarps = 16;                     // 4 per TMEM lane quadrant: one head each in the row mapping
constexpr int kHH = 4 * kHeadsPerItem / kEpiWarps; // heads per thread in the row mapping (1)
constexpr int kTRows = 16 * kTile / (kEpiWarps * 32); // rows per thread in the transform mapping (4)
constexpr int kTStep = kTile / kTRows;             // row stride of the transform mapping (32)
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreads = 64 + kEpiThreads + 32;
constexpr int kTmemCols = 512;
constexpr int kNB = 48;                            // N of the num MMA: 32 KV columns + Ks (-> den) + padding to 16
constexpr int kBmatBytes = kD * kNB * 2;           // one [32 x 48] bf16 operand: columns 0..31 = X_head, column 32 = Ks_head
constexpr int kVecBytes = kTile * 16 * 2;          // a [128 tokens x 16] MN-major operand (ones / per-row weights)
// TMEM columns
constexpr uint32_t kColKV = 0;     // pass A accumulator (dead once the operands are built); pass B: dQ~
constexpr uint32_t kColDQ = 0;
constexpr uint32_t kColDKV = 128;  // pass B accumulator
constexpr uint32_t kColNum = 256;  // pass B: [num | den] per head, 4 x 48 columns
constexpr uint32_t kColDks = 448;  // pass B: Q~^T [w_hi | w_lo] (16 columns)
constexpr uint32_t kColKs = 464;   // pass A: relu(K)^T 1 (16 columns, column 0 = Ks)
// pass C accumulators, double-buffered by tile parity: dV at 0 / 128, dK at 256 / 384 (dKV and dQ~ are dead by then)
constexpr uint32_t kColDV = 0;
constexpr uint32_t kColDK = 256;
constexpr uint32_t kColBuf = 128;

struct BwdSmem {
  uint8_t slots[kStages][kSlotBytes];           // 1024-aligned
  uint8_t bmat[kHeadsPerItem][2][kBmatBytes];   // [head][hi|lo] [32 x 48]: [KV | Ks] in pass B, dKV in pass C
  uint8_t ones[kVecBytes];                      // [128 tokens x 16] MN-major, column 0 = 1: relu(K)^T ones = Ks on the tensor core
  uint8_t wm[kVecBytes];                        // same layout, columns h / 4 + h = hi / lo of (dDen / r) of head h: Q~^T wm = dKs
  float ks[kHeadsPerItem * kD];
  float dks[kHeadsPerItem * kD];
  uint64_t full[kStages], empty[kStages], ready[kStages], out_ready[kStages];
  uint64_t kv_full, b_ready, dkv_full, b2_ready;
  uint64_t num_full, num_empty, dq_full, dq_empty, pair_done, c_full[2], c_empty[2];
  uint32_t tmem_base;
};
constexpr int kBwdSmemBytes = (int)sizeof(BwdSmem) + 1024;
static_assert(kBwdSmemBytes <= 232448, "backward kernel exceeds the 227 KB shared-memory limit");

constexpr uint32_t kIdescTT = make_idesc_bf16(128, 128, 1, 1);  // X^T Y over tokens: A, B MN-major
constexpr uint32_t kIdescMN = make_idesc_bf16(128, 32, 0, 1);   // rows x 32x32 operand read as [k][n]
constexpr uint32_t kIdescNum = make_idesc_bf16(128, kNB, 0, 1); // rows x [32 x 48] operand: numerator and den in one MMA
constexpr uint32_t kIdescVec = make_idesc_bf16(128, 16, 1, 1);  // X^T [tokens x 16] over tokens: column sums on the tensor core
constexpr uint32_t kIdescK = make_idesc_bf16(128, 32, 0, 0);    // rows x transposed 32x32 operand

__device__ __forceinline__ uint4 relu8(uint4 v) {
  const __nv_bfloat162 z = __float2bfloat162_rn(0.f);
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __hmax2_nan(h[i], z);
  return v;
}
__device__ __forceinline__ uint4 pack8(const float f[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// packed-pair helpers for the row epilogues (FFMA2 / FMUL2: one issue slot per two columns; common.cuh)
__device__ __forceinline__ f2 acc2(const uint32_t (&r)[32], int i) {      // columns 2i, 2i+1 of a tcgen05.ld row
  return mk2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
}
__device__ __forceinline__ uint32_t word(const uint4& v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w)); }
// keep the halves of a packed bf16 pair whose mask half is > 0 (the ReLU mask of a gradient: q > 0 / k > 0), zero the rest
__device__ __forceinline__ uint32_t mask_pos(uint32_t val, uint32_t m) {
  const __nv_bfloat162 keep = __hgt2(as_bf162(m), __float2bfloat162_rn(0.f));   // 1.0 / 0.0 per half
  return as_u32(__hmul2(as_bf162(val), keep));
}

// 32x32 fp32 block held as one row (32 values) per lane -> bf16 hi + lo, MN-major no-swizzle core matrices:
// element [k][n] at (n/8)*512 + (k/8)*128 + (k%8)*16 + (n%8)*2.  Read with b_major = MN it is the operand
// X[k][n]; read with b_major = K (LBO 512, SBO 128) the very same bytes are X^T.
__device__ __forceinline__ void store_bmat_hilo(uint8_t* bh, uint8_t* bl, int lane, const uint32_t (&r)[32]) {
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
}

struct BwdMaps {
  CUtensorMap q, k, v, d_o, dq, dk, dv;
};


__global__ void __launch_bounds__(kThreads, 1)
#ifdef YAT_BWD_PROBE   // timing experiment build only (scripts/linattn_bwd_probe.py); never defined for the product library
linattn_bwd_sm100_kernel(const __grid_constant__ BwdMaps M, int B, int N, int groups, int probe_mask) {
#else
linattn_bwd_sm100_kernel(const __grid_constant__ BwdMaps M, int B, int N, int groups) {
#endif
  extern __shared__ uint8_t smem_raw[];
  BwdSmem& S = *reinterpret_cast<BwdSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = (N + kTile - 1) / kTile;
  const int n_items = B * groups;
#ifdef YAT_BWD_PROBE   // timing experiment only (results are wrong): run a subset of the three passes
  const int nA = (probe_mask & 1) ? n_tiles : 0, nB = (probe_mask & 2) ? n_tiles : 0, nC = (probe_mask & 4) ? n_tiles : 0;
#define PROBE_SKIP(bit) (probe_mask & (bit))
#else
#define PROBE_SKIP(bit) false
  const int nA = n_tiles, nB = n_tiles, nC = n_tiles;
#endif

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(smem_u32(&S.full[s]), 1);
      mbar_init(smem_u32(&S.empty[s]), 1);
      mbar_init(smem_u32(&S.ready[s]), kEpiWarps);
      mbar_init(smem_u32(&S.out_ready[s]), kEpiWarps);
    }
    mbar_init(smem_u32(&S.kv_full), 1);
    mbar_init(smem_u32(&S.b_ready), kEpiWarps);
    mbar_init(smem_u32(&S.dkv_full), 1);
    mbar_init(smem_u32(&S.b2_ready), kEpiWarps);
    mbar_init(smem_u32(&S.num_full), 1);
    mbar_init(smem_u32(&S.num_empty), kEpiWarps);
    mbar_init(smem_u32(&S.dq_full), 1);
    mbar_init(smem_u32(&S.dq_empty), kEpiWarps);
    mbar_init(smem_u32(&S.pair_done), 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&S.c_full[i]), 1);
      mbar_init(smem_u32(&S.c_empty[i]), kEpiWarps);
    }
    fence_barrier_init();
    tma_prefetch_desc(&M.q); tma_prefetch_desc(&M.k); tma_prefetch_desc(&M.v); tma_prefetch_desc(&M.d_o);
    tma_prefetch_desc(&M.dq); tma_prefetch_desc(&M.dk); tma_prefetch_desc(&M.dv);
  }
  // operand padding stays zero for the whole kernel: bmat columns 33..47, wm / ones columns 8..15; ones column 0 = 1
  for (int i = threadIdx.x; i < (int)(sizeof(S.bmat) + sizeof(S.ones) + sizeof(S.wm)) / 16; i += kThreads)
    reinterpret_cast<uint4*>(&S.bmat[0][0][0])[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  if (threadIdx.x < kTile)   // element [token][n] of a vector operand at (n/8)*2048 + (token/8)*128 + (token%8)*16 + (n%8)*2
    *reinterpret_cast<uint32_t*>(&S.ones[(threadIdx.x >> 3) * 128 + (threadIdx.x & 7) * 16]) = 0x00003f80u;   // bf16 1.0
  fence_proxy_async();
  if (warp == 1) {
    tmem_alloc(smem_u32(&S.tmem_base), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // One CTA per SM allocating all 512 columns: the only possible base is column 0.  Treating it as the constant it is
  // keeps every TMEM address in the issuer's MMA loop a compile-time immediate (a register-held base costs an
  // ELECT / R2UR.BROADCAST waterfall per tcgen05.mma: ~100 cycles per issue from a single thread).
  if (S.tmem_base != 0) __trap();
  constexpr uint32_t tmem = 0;

  if (warp == 0) {
    // ======================= TMA producer =======================
    // (An L2 prefetch of every tile pair 4 pairs ahead of its ring load - cp.async.bulk.prefetch.tensor - was measured
    // and made every pass SLOWER, 165 -> 188 us: the prefetches queue in the same TMA unit as the loads.)
    if (lane == 0) {
      uint32_t it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int b = item / groups, c0 = (item % groups) * (kHeadsPerItem * kD);
        for (int pass = 0; pass < 3; ++pass) {
          const int np = pass == 0 ? nA : (pass == 1 ? nB : nC);
          for (int t = 0; t < np; ++t) {
#pragma unroll
            for (int w = 0; w < 2; ++w, ++it) {
              const CUtensorMap* tm = pass == 1 ? (w ? &M.d_o : &M.q) : (w ? &M.v : &M.k);
              const uint32_t s = it % kStages, ph = (it / kStages) & 1;
              TRACE(3, 50 + pass);
              mbar_wait(smem_u32(&S.empty[s]), ph ^ 1);
              TRACE(3, 55 + pass);
              const uint32_t full = smem_u32(&S.full[s]);
              mbar_arrive_expect_tx(full, kSlotBytes);
              const uint32_t dst = smem_u32(&S.slots[s][0]);
              tma_load_3d(dst, tm, full, c0, t * kTile, b);
              tma_load_3d(dst + kHalfBytes, tm, full, c0 + 64, t * kTile, b);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================= tcgen05 issuer =======================
    if (lane == 0) {
      uint32_t it = 0, bt = 0, ct = 0, item_iter = 0;
      uint32_t ready_ph = 0;  // per-slot parity of `ready` (V uses of a slot do not complete a phase)
      auto wait_ready = [&](uint32_t s) {
        mbar_wait(smem_u32(&S.ready[s]), (ready_ph >> s) & 1);
        ready_ph ^= 1u << s;
      };
      // rows x (32x32 operand): 4 heads x (hi, lo) x 2 K-steps into 32-column accumulators.  A descriptor is its base
      // plus (byte offset >> 4) as long as the address field does not overflow (shared memory < 256 KB): one 64-bit add
      // of an immediate per MMA instead of re-encoding.
      const uint64_t bmat_mn = make_smem_desc(smem_u32(&S.bmat[0][0][0]), 128, 512, kLayoutNone);   // operand [k][n] as stored
      const uint64_t bmat_k = make_smem_desc(smem_u32(&S.bmat[0][0][0]), 512, 128, kLayoutNone);    // same bytes as the transpose
      auto small_mmas = [&](uint32_t acc_col, uint32_t a_slot, bool transposed, uint32_t idesc_mn = kIdescMN,
                            uint32_t head_cols = kD) {
        const uint64_t a0 = make_smem_desc(a_slot, 16, 1024, kLayoutSw128);
#pragma unroll
        for (int h = 0; h < kHeadsPerItem; ++h) {
          const uint32_t d = tmem + acc_col + h * head_cols;
          const uint32_t aoff = (h >> 1) * kHalfBytes + (h & 1) * 64;  // 32 columns = 64 B inside the 128 B row
#pragma unroll
          for (int hl = 0; hl < (PROBE_SKIP(64) ? 1 : 2); ++hl) {
            const uint32_t boff = (h * 2 + hl) * kBmatBytes;
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
              const uint64_t da = a0 + ((aoff + ks * 32) >> 4);
              if (!transposed) {
                // MN-major: 8-k group stride (LBO) 128 B, 8-n group stride (SBO) 512 B; K step of 16 = 256 B
                umma_bf16(d, da, bmat_mn + ((boff + ks * 256) >> 4), idesc_mn, (hl | ks) != 0);
              } else {
                // K-major view: 8-k chunk stride (LBO) 512 B, 8-n group stride (SBO) 128 B; K step of 16 = 1024 B
                umma_bf16(d, da, bmat_k + ((boff + ks * 1024) >> 4), kIdescK, (hl | ks) != 0);
              }
            }
          }
        }
      };
      // [128 x 16] += X^T vec over the 128 tokens: column sums / weighted column sums of a tile on the tensor core
      // (vec: MN-major no-swizzle [tokens][16], 8-token group stride 128 B, 8-column group stride 2048 B)
      auto vec_mmas = [&](uint32_t acc_col, uint32_t x_slot, uint32_t vec_addr, bool first) {
        const uint64_t x0 = make_smem_desc(x_slot, kHalfBytes, 1024, kLayoutSw128);
        const uint64_t v0 = make_smem_desc(vec_addr, 128, 2048, kLayoutNone);
#pragma unroll
        for (int ks = 0; ks < kTile / 16; ++ks)
          umma_bf16(tmem + acc_col, x0 + ((ks * 2048) >> 4), v0 + ((ks * 256) >> 4), kIdescVec, !(first && ks == 0));
      };
      // [128 x 128] += X^T Y over the 128 tokens of a tile pair (8 K-steps of 16 tokens = 2048 B each)
      auto big_mmas = [&](uint32_t acc_col, uint32_t x_slot, uint32_t y_slot, bool first) {
        const uint64_t x0 = make_smem_desc(x_slot, kHalfBytes, 1024, kLayoutSw128);
        const uint64_t y0 = make_smem_desc(y_slot, kHalfBytes, 1024, kLayoutSw128);
#pragma unroll
        for (int ks = 0; ks < kTile / 16; ++ks)
          umma_bf16(tmem + acc_col, x0 + ((ks * 2048) >> 4), y0 + ((ks * 2048) >> 4), kIdescTT, !(first && ks == 0));
      };
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++item_iter) {
        // TMEM columns [0,128) held dV of the previous item: its epilogue has run, because the first MMA below waits for
        // ready[] of THIS item's pass A, which every epilogue warp signals only after finishing the previous item.
        // ---- pass A: D_KV += relu(K)^T V
        for (int t = 0; t < nA; ++t, it += 2) {
          const uint32_t sk = it % kStages, sv = (it + 1) % kStages, pv = ((it + 1) / kStages) & 1;
          wait_ready(sk);
          mbar_wait(smem_u32(&S.full[sv]), pv);
          tc_fence_after();
          big_mmas(kColKV, smem_u32(&S.slots[sk][0]), smem_u32(&S.slots[sv][0]), t == 0);
          vec_mmas(kColKs, smem_u32(&S.slots[sk][0]), smem_u32(&S.ones[0]), t == 0);      // Ks = relu(K)^T 1
          tc_commit(smem_u32(&S.empty[sk]));
          tc_commit(smem_u32(&S.empty[sv]));
        }
        tc_commit(smem_u32(&S.kv_full));
        mbar_wait(smem_u32(&S.b_ready), item_iter & 1);
        tc_fence_after();
        // ---- pass B.  Tensor-pipe order per tile: dQ~(t), num(t+1), dKV(t) - the epilogue warps transform tile t+1
        // (E1) before they finish tile t, so num(t+1) is in TMEM by the time E2(t+1) wants it.
        if (nB > 0) {
          const uint32_t sq0 = it % kStages;
          wait_ready(sq0);                                       // relu(Q) in place, den written
          mbar_wait(smem_u32(&S.num_empty), (bt & 1) ^ 1);
          tc_fence_after();
          small_mmas(kColNum, smem_u32(&S.slots[sq0][0]), false, kIdescNum, kNB);   // [num | den] = relu(Q) [KV | Ks]
          tc_commit(smem_u32(&S.num_full));
        }
        for (int t = 0; t < nB; ++t, it += 2, ++bt) {
          const uint32_t sq = it % kStages, sd = (it + 1) % kStages;
          const uint32_t qa = smem_u32(&S.slots[sq][0]), da_ = smem_u32(&S.slots[sd][0]);
          TRACE(1, 30);
          // dQ~ = dO KV^T needs only the raw dO tile and a drained dQ accumulator: it is issued while the epilogue warps
          // are still in E2 of this tile, so dq_full is up by the time E3b asks for it
          mbar_wait(smem_u32(&S.full[sd]), ((it + 1) / kStages) & 1);
          mbar_wait(smem_u32(&S.dq_empty), (bt & 1) ^ 1);
          tc_fence_after();
          TRACE(1, 31);
          small_mmas(kColDQ, da_, true);
          tc_commit(smem_u32(&S.dq_full));
          TRACE(1, 32);
          wait_ready(sd);                                        // Q~ written in place (Q slot)
          tc_fence_after();
          TRACE(1, 33);
          // the tile's token-axis contractions first (pair_done frees the dO / Q~ slots and lets E3b write dQ), then the
          // numerator of the next tile: E2(t+1) is a whole epilogue away, E3b(t) is waiting now
          big_mmas(kColDKV, qa, da_, t == 0);                    // D_dKV += Q~^T dO
          vec_mmas(kColDks, qa, smem_u32(&S.wm[0]), t == 0);     // D_dKs += Q~^T [w_hi | w_lo]
          tc_commit(smem_u32(&S.pair_done));
          if (t + 1 < nB) {
            const uint32_t sq1 = (it + 2) % kStages;
            wait_ready(sq1);
            mbar_wait(smem_u32(&S.num_empty), ((bt + 1) & 1) ^ 1);
            tc_fence_after();
            small_mmas(kColNum, smem_u32(&S.slots[sq1][0]), false, kIdescNum, kNB);
            tc_commit(smem_u32(&S.num_full));
          }
          TRACE(1, 34);
        }
        tc_commit(smem_u32(&S.dkv_full));
        mbar_wait(smem_u32(&S.b2_ready), item_iter & 1);
        tc_fence_after();
        // ---- pass C: dV = relu(K) dKV, dK~ = V dKV^T
        for (int t = 0; t < nC; ++t, it += 2, ++ct) {
          const uint32_t sk = it % kStages, sv = (it + 1) % kStages, pv = ((it + 1) / kStages) & 1;
          const uint32_t buf = ct & 1;
          TRACE(1, 70);
          wait_ready(sk);
          mbar_wait(smem_u32(&S.full[sv]), pv);
          TRACE(1, 71);
          mbar_wait(smem_u32(&S.c_empty[buf]), ((ct >> 1) & 1) ^ 1);
          tc_fence_after();
          TRACE(1, 72);
          small_mmas(kColDV + buf * kColBuf, smem_u32(&S.slots[sk][0]), false);
          small_mmas(kColDK + buf * kColBuf, smem_u32(&S.slots[sv][0]), true);
          tc_commit(smem_u32(&S.c_full[buf]));
          TRACE(1, 73);
        }
      }
    }
  } else if (warp == 2 + kEpiWarps) {
    // ======================= TMA store + slot release =======================
    if (lane == 0) {
      uint32_t it = 0, out_ph = 0;
      auto wait_out = [&](uint32_t s) {
        mbar_wait(smem_u32(&S.out_ready[s]), (out_ph >> s) & 1);
        out_ph ^= 1u << s;
      };
      auto store = [&](const CUtensorMap* tm, uint32_t s, int c0, int row0, int b) {
        const uint32_t src = smem_u32(&S.slots[s][0]);
        tma_store_3d(tm, src, c0, row0, b);
        tma_store_3d(tm, src + kHalfBytes, c0 + 64, row0, b);
      };
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int b = item / groups, c0 = (item % groups) * (kHeadsPerItem * kD);
        it += 2 * nA;  // pass A: slots are released by the issuer's commits
        for (int t = 0; t < nB; ++t, it += 2) {
          const uint32_t sq = it % kStages, sd = (it + 1) % kStages;
          TRACE(2, 40);
          wait_out(sd);                                  // dQ sits in the dO slot; nobody reads the Q slot any more
          TRACE(2, 41);
          mbar_arrive(smem_u32(&S.empty[sq]));
          store(&M.dq, sd, c0, t * kTile, b);
          tma_store_commit();
          tma_store_wait_read<0>();
          mbar_arrive(smem_u32(&S.empty[sd]));
          TRACE(2, 42);
        }
        for (int t = 0; t < nC; ++t, it += 2) {
          const uint32_t sk = it % kStages, sv = (it + 1) % kStages;
          TRACE(2, 80);
          wait_out(sk);
          TRACE(2, 81);
          store(&M.dk, sk, c0, t * kTile, b);
          tma_store_commit();
          tma_store_wait_read<0>();                      // dV's epilogue is still running: release the K slot now
          mbar_arrive(smem_u32(&S.empty[sk]));
          wait_out(sv);
          store(&M.dv, sv, c0, t * kTile, b);
          tma_store_commit();
          tma_store_wait_read<0>();
          mbar_arrive(smem_u32(&S.empty[sv]));
          TRACE(2, 82);
        }
      }
      tma_store_wait_all<0>();
    }
  } else {
    // ======================= transform + epilogue warps (2..9) =======================
    const int te = threadIdx.x - 64;          // 0..kEpiThreads-1
    const int we = te >> 5;                   // 0..kEpiWarps-1
    const int cc = te & 15;                   // transform mapping: 16-byte chunk column of the 256-byte item row
    const int rbase = te >> 4;                //                    rows rbase + kTStep i
    const int half = cc >> 3, chunk = cc & 7;
    const int quad = warp & 3;                // TMEM lane quadrant this warp may access
    const int colsel = we >> 2;               // row mapping: heads colsel * kHH .. + kHH - 1 are finished by this warp
    const int row = quad * 32 + lane;         //              token row in the tile (TMEM lane)
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    // head h of the row: 64-column half h >> 1, 16-byte chunks (h & 1) * 4 + g (swizzled by the row)
    const int h0 = colsel * kHH;
    const int row_off = (h0 >> 1) * kHalfBytes + row * 128;
    const int ch0 = (h0 & 1) * 4;
    uint32_t it = 0, bt = 0, ct = 0, item_iter = 0;

    // diagonal 32x32 blocks of a [128x128] accumulator -> hi/lo operands in S.bmat (warps with we < 4: head = quad)
    // The column sums that go with it come off the tensor core too (this thread's TMEM lane = column quad*32 + lane):
    // with_ks: Ks = relu(K)^T 1 -> S.ks and column 32 of the operand (so the num MMA also yields den);
    // else:    dKs = Q~^T [w_hi | w_lo] -> S.dks.
    auto build_bmat = [&](uint32_t col, bool with_ks) {
      if (we < 4) {
        uint32_t r[32];
        tmem_ld32(tmem + lane_addr + col + quad * kD, r);   // X_head[k = lane][n = 0..31]
        uint32_t v8[8];
        tmem_ld8(tmem + lane_addr + (with_ks ? kColKs : kColDks), v8);
        tmem_ld_wait();
        store_bmat_hilo(&S.bmat[quad][0][0], &S.bmat[quad][1][0], lane, r);
        if (with_ks) {
          const float ks = __uint_as_float(v8[0]);
          S.ks[quad * kD + lane] = ks;
          const uint32_t khi = pack_bf16(ks, 0.f);
          const uint32_t klo = pack_bf16(ks - bf16_lo(khi), 0.f);
          const int off = 4 * 512 + (lane >> 3) * 128 + (lane & 7) * 16;   // [k = lane][n = 32]
          *reinterpret_cast<uint4*>(&S.bmat[quad][0][off]) = make_uint4(khi, 0, 0, 0);
          *reinterpret_cast<uint4*>(&S.bmat[quad][1][off]) = make_uint4(klo, 0, 0, 0);
        } else {
          float hi = 0.f, lo = 0.f;
#pragma unroll
          for (int j = 0; j < kHeadsPerItem; ++j) {                        // select column `quad` without dynamic indexing
            hi = j == quad ? __uint_as_float(v8[j]) : hi;
            lo = j == quad ? __uint_as_float(v8[kHeadsPerItem + j]) : lo;
          }
          S.dks[quad * kD + lane] = hi + lo;
        }
        fence_proxy_async();
      }
      tc_fence_before();
      named_bar_sync(1, kEpiThreads);
    };

    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++item_iter) {
      // ---------------- pass A: relu(K) in place, Ks ----------------
      for (int t = 0; t < nA; ++t, it += 2) {
        const uint32_t sk = it % kStages, pk = (it / kStages) & 1;
        mbar_wait(smem_u32(&S.full[sk]), pk);
        uint8_t* base = &S.slots[sk][half * kHalfBytes];
#pragma unroll
        for (int i = 0; i < kTRows; ++i) {
          const int r = rbase + kTStep * i;
          uint4* p = reinterpret_cast<uint4*>(base + r * 128 + ((chunk ^ (r & 7)) << 4));
          *p = relu8(*p);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.ready[sk]));
        mbar_wait(smem_u32(&S.full[(it + 1) % kStages]), ((it + 1) / kStages) & 1);  // observe the V phase too
      }
      mbar_wait(smem_u32(&S.kv_full), item_iter & 1);
      tc_fence_after();
      build_bmat(kColKV, true);                 // barrier inside: S.ks visible, D_KV drained, [KV | Ks] operands written
      if (lane == 0) mbar_arrive(smem_u32(&S.b_ready));
      // ---------------- pass B ----------------
      // Software-pipelined by one tile: E1(t+1) (relu(Q), den) runs BEFORE E2(t) / E3(t), so the num MMA of tile t+1
      // and its mbarrier round trip overlap this tile's epilogue instead of sitting on the critical path.
      // E1: relu(Q) in place.  den = relu(Q).Ks comes out of the num MMA (column 32 of the [KV | Ks] operand) and
      // dKs = sum_n Q~_n (dDen_n / r_n) out of a 16-column MMA on the weights E2 leaves in S.wm: no CUDA-core dot
      // products, no cross-thread traffic through shared memory and no CTA-wide barrier are left in this pass.
      auto e1 = [&](uint32_t jt) {
        const uint32_t sq = jt % kStages, pq = (jt / kStages) & 1;
        mbar_wait(smem_u32(&S.full[sq]), pq);
        uint8_t* base = &S.slots[sq][half * kHalfBytes];
#pragma unroll
        for (int i = 0; i < kTRows; ++i) {
          const int r = rbase + kTStep * i;
          uint4* p = reinterpret_cast<uint4*>(base + r * 128 + ((chunk ^ (r & 7)) << 4));
          *p = relu8(*p);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.ready[sq]));
      };
      if (nB > 0) e1(it);
      for (int t = 0; t < nB; ++t, it += 2, ++bt) {
        const uint32_t sq = it % kStages;
        const uint32_t sd = (it + 1) % kStages, pd = ((it + 1) / kStages) & 1;
        if (te == 0) TRACE(0, 10);
        if (t + 1 < nB) e1(it + 2);
        if (te == 0) TRACE(0, 11);

        // E2: num -> r, dDen; Q~ = bf16(r * relu(Q)) in place
        mbar_wait(smem_u32(&S.full[sd]), pd);
        if (te == 0) TRACE(0, 13);
        mbar_wait(smem_u32(&S.num_full), bt & 1);
        tc_fence_after();
        if (te == 0) TRACE(0, 14);
        float rr[kHH], dd[kHH];
        uint8_t* qrow = &S.slots[sq][row_off];
        const uint8_t* drow = &S.slots[sd][row_off];
        #pragma unroll
        for (int hh = 0; hh < kHH; ++hh) rr[hh] = dd[hh] = 0.f;
        if (!PROBE_SKIP(16))
#pragma unroll
        for (int hh = 0; hh < kHH; ++hh) {
          const int h = h0 + hh;
          uint32_t acc[32], dn;
          tmem_ld32(tmem + lane_addr + kColNum + h * kNB, acc);
          tmem_ld1(tmem + lane_addr + kColNum + h * kNB + kD, dn);       // den = relu(Q_n).Ks
          tmem_ld_wait();
          f2 s2 = bc2(0.f);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint4 dv = *reinterpret_cast<const uint4*>(drow + (((ch0 + hh * 4 + g) ^ (row & 7)) << 4));
#pragma unroll
            for (int j = 0; j < 4; ++j) s2 = fma2(unpack2(word(dv, j)), acc2(acc, g * 4 + j), s2);
          }
          float s, s_hi;
          un2(s2, s, s_hi);
          s += s_hi;
          const float r = 1.0f / (__uint_as_float(dn) + 1e-15f);
          rr[hh] = r;
          dd[hh] = -s * r * r;
          {  // dDen / r as bf16 hi + lo, columns h and 4 + h of the [tokens x 16] operand of the dKs MMA
            const float w = -s * r;
            const __nv_bfloat16 whi = __float2bfloat16_rn(w);
            const __nv_bfloat16 wlo = __float2bfloat16_rn(w - __bfloat162float(whi));
            __nv_bfloat16* wr = reinterpret_cast<__nv_bfloat16*>(&S.wm[(row >> 3) * 128 + (row & 7) * 16]);
            wr[h] = whi;
            wr[kHeadsPerItem + h] = wlo;
          }
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int o = ((ch0 + hh * 4 + g) ^ (row & 7)) << 4;
            const uint4 qv = *reinterpret_cast<const uint4*>(qrow + o);
            const f2 r2 = bc2(r);
            *reinterpret_cast<uint4*>(qrow + o) =
                make_uint4(pack2(mul2(unpack2(qv.x), r2)), pack2(mul2(unpack2(qv.y), r2)), pack2(mul2(unpack2(qv.z), r2)),
                           pack2(mul2(unpack2(qv.w), r2)));
          }
        }
        tc_fence_before();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(smem_u32(&S.num_empty));
          mbar_arrive(smem_u32(&S.ready[sd]));
        }
        if (te == 0) TRACE(0, 15);

        // E3b: dQ = (r * dQ~ + dDen * Ks) * [q > 0]
        if (te == 0) TRACE(0, 17);
        mbar_wait(smem_u32(&S.dq_full), bt & 1);
        tc_fence_after();
        if (te == 0) TRACE(0, 18);
        uint4 outv[kHH][4];
#pragma unroll
        for (int hh = 0; hh < kHH; ++hh)
#pragma unroll
          for (int g = 0; g < 4; ++g) outv[hh][g] = make_uint4(0, 0, 0, 0);
        if (!PROBE_SKIP(32))
#pragma unroll
        for (int hh = 0; hh < kHH; ++hh) {
          const int h = h0 + hh;
          uint32_t acc[32];
          tmem_ld32(tmem + lane_addr + kColDQ + h * kD, acc);
          tmem_ld_wait();
          const f2 r2 = bc2(rr[hh]), d2 = bc2(dd[hh]);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint4 mv = *reinterpret_cast<const uint4*>(qrow + (((ch0 + hh * 4 + g) ^ (row & 7)) << 4));   // Q~ >= 0: the mask
            const float4 k0 = *reinterpret_cast<const float4*>(&S.ks[h * kD + g * 8]);       // broadcast 128-bit loads
            const float4 k1 = *reinterpret_cast<const float4*>(&S.ks[h * kD + g * 8 + 4]);
            const f2 kk[4] = {mk2(k0.x, k0.y), mk2(k0.z, k0.w), mk2(k1.x, k1.y), mk2(k1.z, k1.w)};
            uint32_t w4[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
              w4[j] = mask_pos(pack2(fma2(r2, acc2(acc, g * 4 + j), mul2(d2, kk[j]))), word(mv, j));
            outv[hh][g] = make_uint4(w4[0], w4[1], w4[2], w4[3]);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.dq_empty));
        if (te == 0) TRACE(0, 19);
        mbar_wait(smem_u32(&S.pair_done), bt & 1);   // the MMAs have finished reading the dO / Q~ tiles
        if (te == 0) TRACE(0, 20);
        uint8_t* orow = &S.slots[sd][row_off];
#pragma unroll
        for (int hh = 0; hh < kHH; ++hh)
#pragma unroll
          for (int g = 0; g < 4; ++g)
            *reinterpret_cast<uint4*>(orow + (((ch0 + hh * 4 + g) ^ (row & 7)) << 4)) = outv[hh][g];
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.out_ready[sd]));
      }
      mbar_wait(smem_u32(&S.dkv_full), item_iter & 1);
      tc_fence_after();
      build_bmat(kColDKV, false);               // barrier inside: S.dks visible, D_dKV / D_dKs drained, operands rewritten
      if (lane == 0) mbar_arrive(smem_u32(&S.b2_ready));
      // ---------------- pass C ----------------
      // Same skew: relu(K) of tile t+1 is done before the epilogue of tile t, and the dV / dK accumulators are
      // double-buffered in TMEM (tile parity), so the issuer runs the MMAs of tile t+1 during this tile's epilogue.
      auto relu_k = [&](uint32_t jt) {
        const uint32_t sk = jt % kStages, pk = (jt / kStages) & 1;
        mbar_wait(smem_u32(&S.full[sk]), pk);
        uint8_t* base = &S.slots[sk][half * kHalfBytes];
#pragma unroll
        for (int i = 0; i < kTRows; ++i) {
          const int r = rbase + kTStep * i;
          uint4* p = reinterpret_cast<uint4*>(base + r * 128 + ((chunk ^ (r & 7)) << 4));
          *p = relu8(*p);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.ready[sk]));
      };
      if (nC > 0) relu_k(it);
      for (int t = 0; t < nC; ++t, it += 2, ++ct) {
        const uint32_t sk = it % kStages, sv = (it + 1) % kStages;
        const uint32_t buf = ct & 1;
        if (te == 0) TRACE(0, 60);
        if (t + 1 < nC) relu_k(it + 2);
        if (te == 0) TRACE(0, 61);
        mbar_wait(smem_u32(&S.full[sv]), ((it + 1) / kStages) & 1);   // observe the V phase (complete: the MMA read it)
        mbar_wait(smem_u32(&S.c_full[buf]), (ct >> 1) & 1);   // both MMAs done: K and V tiles may be overwritten
        tc_fence_after();
        if (te == 0) TRACE(0, 62);
        // dK = (V dKV^T + dKs) * [k > 0]  -> K slot
        uint8_t* krow = &S.slots[sk][row_off];
#pragma unroll
        for (int hh = 0; hh < kHH; ++hh) {
          const int h = h0 + hh;
          uint32_t acc[32];
          tmem_ld32(tmem + lane_addr + kColDK + buf * kColBuf + h * kD, acc);
          tmem_ld_wait();
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4* p = reinterpret_cast<uint4*>(krow + (((ch0 + hh * 4 + g) ^ (row & 7)) << 4));
            const uint4 mv = *p;                                                             // relu(k) >= 0: the mask
            const float4 k0 = *reinterpret_cast<const float4*>(&S.dks[h * kD + g * 8]);      // broadcast 128-bit loads
            const float4 k1 = *reinterpret_cast<const float4*>(&S.dks[h * kD + g * 8 + 4]);
            const f2 kk[4] = {mk2(k0.x, k0.y), mk2(k0.z, k0.w), mk2(k1.x, k1.y), mk2(k1.z, k1.w)};
            uint32_t w4[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) w4[j] = mask_pos(pack2(add2(acc2(acc, g * 4 + j), kk[j])), word(mv, j));
            *p = make_uint4(w4[0], w4[1], w4[2], w4[3]);
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&S.out_ready[sk]));
        if (te == 0) TRACE(0, 63);
        // dV = relu(K) dKV -> V slot
        uint8_t* vrow = &S.slots[sv][row_off];
#pragma unroll
        for (int hh = 0; hh < kHH; ++hh) {
          const int h = h0 + hh;
          uint32_t acc[32];
          tmem_ld32(tmem + lane_addr + kColDV + buf * kColBuf + h * kD, acc);
          tmem_ld_wait();
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            float o8[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o8[j] = __uint_as_float(acc[g * 8 + j]);
            *reinterpret_cast<uint4*>(vrow + (((ch0 + hh * 4 + g) ^ (row & 7)) << 4)) = pack8(o8);
          }
        }
        tc_fence_before();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(smem_u32(&S.c_empty[buf]));
          mbar_arrive(smem_u32(&S.out_ready[sv]));
        }
        if (te == 0) TRACE(0, 64);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

}  // namespace
}  // namespace yat


YAT_DEFINE_TRACE_API(yat_bwd)

extern "C" int yat_linattn_bwd(const void* q, const void* k, const void* v, const void* dout, int B, int N, int H,
                               int D, void* dq, void* dk, void* dv, void* stream) {
  return yat_linattn_bwd_ld(q, k, v, (int64_t)H * D, dout, B, N, H, D, dq, dk, dv, (int64_t)H * D, stream);
}

extern "C" int yat_linattn_bwd_ld(const void* q, const void* k, const void* v, int64_t ld, const void* dout, int B,
                                  int N, int H, int D, void* dq, void* dk, void* dv, int64_t ld_grad, void* stream) {
  using namespace yat;
  YAT_REQUIRE(ld >= (int64_t)H * D && ld % 8 == 0 && ld_grad >= (int64_t)H * D && ld_grad % 8 == 0, YAT_ERR_BAD_ARG,
              "yat_linattn_bwd: row pitches %lld / %lld must be >= H*D and multiples of 8", (long long)ld, (long long)ld_grad);
  YAT_REQUIRE(q && k && v && dout && dq && dk && dv, YAT_ERR_BAD_ARG, "yat_linattn_bwd: null pointer argument");
  YAT_REQUIRE(B > 0 && N > 0 && H > 0, YAT_ERR_BAD_ARG, "yat_linattn_bwd: non-positive dimension");
  YAT_REQUIRE(D == kD, YAT_ERR_UNSUPPORTED, "yat_linattn_bwd: head_dim %d not supported (SANA uses 32)", D);
  YAT_REQUIRE((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)dout | (uintptr_t)dq | (uintptr_t)dk |
                (uintptr_t)dv) & 15u) == 0,
              YAT_ERR_BAD_ARG, "yat_linattn_bwd: tensors must be 16-byte aligned");
  BwdMaps maps;
  const uint64_t C = (uint64_t)H * D;
  const void* ptrs[7] = {q, k, v, dout, dq, dk, dv};
  CUtensorMap* tms[7] = {&maps.q, &maps.k, &maps.v, &maps.d_o, &maps.dq, &maps.dk, &maps.dv};
  for (int i = 0; i < 7; ++i) {
    const uint64_t pitch = i < 3 ? (uint64_t)ld : (i == 3 ? C : (uint64_t)ld_grad);   // q k v | dO | dq dk dv
    const int rc = make_tmap_3d_bf16(tms[i], ptrs[i], C, N, B, 64, kTile, pitch);
    if (rc != YAT_OK) return rc;
  }
  static bool attr_set = false;
  if (!attr_set) {
    const cudaError_t e =
        cudaFuncSetAttribute(linattn_bwd_sm100_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmemBytes);
    YAT_REQUIRE(e == cudaSuccess, YAT_ERR_CUDA, "yat_linattn_bwd: cannot reserve %d B of shared memory (%s)",
                kBwdSmemBytes, cudaGetErrorString(e));
    attr_set = true;
  }
  const int groups = (H + kHeadsPerItem - 1) / kHeadsPerItem;
  const int items = B * groups;
  const int grid = items < kNumSMs ? items : kNumSMs;
#ifdef YAT_BWD_PROBE
  int probe_mask = 7;
  if (const char* e = getenv("YAT_BWD_MASK")) probe_mask = atoi(e);
  linattn_bwd_sm100_kernel<<<grid, kThreads, kBwdSmemBytes, (cudaStream_t)stream>>>(maps, B, N, groups, probe_mask);
#else
  linattn_bwd_sm100_kernel<<<grid, kThreads, kBwdSmemBytes, (cudaStream_t)stream>>>(maps, B, N, groups);
#endif
  return check_launch("linattn_bwd_sm100");
}
