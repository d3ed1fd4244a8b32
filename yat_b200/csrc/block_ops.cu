// Fused neighbours of the attention / feed-forward sub-layers of a SANA block:
//
//   ln_modulate     y = LayerNorm(x) * (1 + scale_b) + shift_b          (norm1/norm2 + adaLN-single modulate,
//                                                                        utils/patch_sana_attention_layers.py:84-90,104-105;
//                                                                        SanaModulatedNorm, patched_sana_transformer.py norm_out)
//   gate_residual   out = res + gate_b * y                              (:95 and :113)
//
// x, y, res, out are [B, N, C] bf16 token-major; scale/shift/gate are per-sample rows [B, C] bf16 (a
// batch stride is passed because they are chunks of one [B, 6, C] tensor).  The reference runs these
// as LayerNorm + 2 broadcast elementwise launches (+ their autograd mirror) per use — 4 passes over
// [16,1024,2240] forward and ~8 backward, most of them on the slow non-vectorised broadcast path;
// here each direction is one pass (plus a column-sum pass for the per-sample gradients).
// All arithmetic in fp32, one rounding to bf16 on the way out.  HBM-bound.
#include <cstdlib>

#include "common.cuh"
#include "sm100.cuh"

namespace yat {
namespace {
using namespace sm100;

constexpr int kWarpsPerBlock = 8;
constexpr int kThreads = kWarpsPerBlock * 32;

__device__ __forceinline__ void unpack8(const uint4& v, float f[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float f[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}
__device__ __forceinline__ uint4 ld_cached(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// ------------------------------------------------------------------------------------------------
// ln_modulate forward: one warp per token row, the row lives in registers (MAXV 128-bit vectors per lane)
// ------------------------------------------------------------------------------------------------
// ADD: the row is first formed as bf16(x + x2) (the attn2 residual add `attn_output + hidden_states`,
// utils/patch_sana_attention_layers.py:103, rounded to bf16 like the reference's tensor add), written to xsum, and
// normalised from the rounded values: one pass (8 C bytes per token) instead of an add pass plus a norm pass (10 C).
template <int MAXV, bool ADD>
__global__ void __launch_bounds__(kThreads) ln_modulate_fwd_kernel(
    const uint4* __restrict__ x, const uint4* __restrict__ x2, const uint4* __restrict__ scale,
    const uint4* __restrict__ shift, int64_t mod_stride_vec, int rows, int N, int nvec, float inv_c, float eps,
    uint4* __restrict__ xsum, uint4* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int b = row / N;
  const uint4* xr = x + (int64_t)row * nvec;
  uint4 v[MAXV];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) v[i] = ld_stream(xr + c);
  }
  if (ADD) {
    const uint4* ar = x2 + (int64_t)row * nvec;
    uint4* sr = xsum + (int64_t)row * nvec;
    uint4 a[MAXV];
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) a[i] = ld_stream(ar + c);
    }
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        v[i].x = pack2(add2(unpack2(v[i].x), unpack2(a[i].x)));
        v[i].y = pack2(add2(unpack2(v[i].y), unpack2(a[i].y)));
        v[i].z = pack2(add2(unpack2(v[i].z), unpack2(a[i].z)));
        v[i].w = pack2(add2(unpack2(v[i].w), unpack2(a[i].w)));
        sr[c] = v[i];                                   // re-read by the residual path: default caching
      }
    }
  }
  // packed fp32 pairs (FFMA2): the row statistics and the modulate take half the issue slots of the scalar form
  f2 s2 = bc2(0.f);
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    if (lane + 32 * i < nvec) {
      const uint32_t w[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) s2 = add2(s2, unpack2(w[j]));
    }
  }
  float slo, shi;
  un2(s2, slo, shi);
  const float mean = warp_sum(slo + shi) * inv_c;
  const f2 nmean = bc2(-mean);
  f2 q2 = bc2(0.f);
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    if (lane + 32 * i < nvec) {
      const uint32_t w[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const f2 d = add2(unpack2(w[j]), nmean);
        q2 = fma2(d, d, q2);
      }
    }
  }
  un2(q2, slo, shi);
  const float rstd = rsqrtf(warp_sum(slo + shi) * inv_c + eps);
  if (lane == 0 && mean_out) {
    mean_out[row] = mean;
    rstd_out[row] = rstd;
  }
  const uint4* sc = scale + (int64_t)b * mod_stride_vec;
  const uint4* sh = shift + (int64_t)b * mod_stride_vec;
  uint4* yr = y + (int64_t)row * nvec;
  const f2 rs = bc2(rstd), one = bc2(1.0f);
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      const uint4 av = ld_cached(sc + c), dv = ld_cached(sh + c);
      const uint32_t w[4] = {v[i].x, v[i].y, v[i].z, v[i].w}, aw[4] = {av.x, av.y, av.z, av.w},
                     dw[4] = {dv.x, dv.y, dv.z, dv.w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)      // ((x - mean) * rstd) * (1 + scale) + shift, same operation order as before
        o[j] = pack2(fma2(mul2(add2(unpack2(w[j]), nmean), rs), add2(unpack2(aw[j]), one), unpack2(dw[j])));
      st_stream(yr + c, make_uint4(o[0], o[1], o[2], o[3]));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// ln_modulate backward, column-ownership form over a shared-memory row ring (the one launched).
//
// A CTA = (row split, sample), one CTA per SM.  Its rows arrive four at a time through bulk asynchronous copies
// (cp.async.bulk global -> shared, completion counted on an mbarrier): the four rows of x, of dy and of the residual
// gradient are each ONE contiguous 4 * 2 C-byte copy issued by a single thread, so the bytes in flight per SM
// (two to three 12 C-row groups, 107-161 KB at C = 2240) do not depend on registers or on how many warps are waiting.
// A thread owns ONE 4-column vector (8 bytes) for the whole CTA, so (1 + scale) is unpacked once into registers and
// the dscale / dshift column sums are plain register accumulators written out at the end (no shared-memory image).
// The 8 partial row sums (4 rows x {sum g, sum g*xhat}) are reduced in the warp with a halving exchange (9 shuffles),
// across the warps through shared memory (lane i of every warp adds value i over the warps in index order, then the
// eight totals are broadcast by shuffle -> identical statistics everywhere, deterministic), ONE barrier per 4 rows.
// A thread copies its 12 vectors of the group from the ring into registers BEFORE that barrier, so the barrier also
// proves that the CTA is done with the slot: it is refilled at once with the group kSlots iterations ahead.
// x, dy and dres are read once; fp32 pairs (FFMA2).
// ------------------------------------------------------------------------------------------------
constexpr int kLnRows = 4;
constexpr int kLnMaxWarps = 32;   // C <= 4096 -> C / 4 <= 1024 threads
constexpr int kLnMaxSlots = 4;

__device__ __forceinline__ void bulk_load(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_dst),
               "l"(gsrc), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds32f(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}

// shared memory: ring [slots][3 tensors][kLnRows rows][C] bf16 | red [2][warps][8] fp32 | barriers
inline int ln_bwd_slots(int C) {
  const int64_t slot = 3ll * kLnRows * C * 2;
  int s = (int)((232448 - 128 - 2 * kLnMaxWarps * 2 * kLnRows * 4 - kLnMaxSlots * 8) / slot);
  return s > kLnMaxSlots ? kLnMaxSlots : s;
}
inline int64_t ln_bwd_smem_bytes(int slots, int C) {
  return 128 + (int64_t)slots * 3 * kLnRows * C * 2 + 2 * kLnMaxWarps * 2 * kLnRows * 4 + kLnMaxSlots * 8;
}

template <int MAXT>
__global__ void __launch_bounds__(MAXT, 1) ln_modulate_bwd_ring_kernel(
    const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ scale,
    int64_t mod_stride, const float* __restrict__ mean_in, const float* __restrict__ rstd_in, int N, int C, int slots,
    int rows_per_split, float inv_c, const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
    float* __restrict__ part0, float* __restrict__ part1) {
  // dres (may be null): gradient arriving on the residual branch of the same tensor; dx = dres + dx_ln is formed in
  // fp32 and rounded once (autograd would round dx_ln, then add in bf16 in a separate 6 C-byte pass)
  extern __shared__ uint8_t ln_smem_raw[];
  const uint32_t row_bytes = (uint32_t)C * 2u, tensor_bytes = kLnRows * row_bytes, slot_bytes = 3u * tensor_bytes;
  const uint32_t ring0 = (smem_u32(ln_smem_raw) + 127u) & ~127u;
  const uint32_t red0 = ring0 + slots * slot_bytes;
  const uint32_t bar0 = red0 + 2 * kLnMaxWarps * 2 * kLnRows * 4;
  const int c = threadIdx.x, lane = c & 31, warp = c >> 5, nwarps = blockDim.x >> 5;
  const int split = blockIdx.x, b = blockIdx.y;
  const bool live = 4 * c < C;
  const int r0 = split * rows_per_split;
  const int r1 = min(N, r0 + rows_per_split);
  const int n_groups = (r1 - r0 + kLnRows - 1) / kLnRows;

  if (c == 0) {
    for (int i = 0; i < slots; ++i) mbar_init(bar0 + i * 8, 1);
    fence_barrier_init();
  }
  __syncthreads();

  auto issue = [&](int g, int slot) {                        // row group g of this CTA -> ring slot (one thread)
    const int rb = r0 + g * kLnRows;
    const uint32_t bytes = (uint32_t)min(kLnRows, r1 - rb) * row_bytes;
    const int64_t off = ((int64_t)b * N + rb) * C;
    const uint32_t dst = ring0 + slot * slot_bytes, bar = bar0 + slot * 8;
    mbar_arrive_expect_tx(bar, dres != nullptr ? 3u * bytes : 2u * bytes);
    bulk_load(dst, x + off, bytes, bar);
    bulk_load(dst + tensor_bytes, dy + off, bytes, bar);
    if (dres != nullptr) bulk_load(dst + 2u * tensor_bytes, dres + off, bytes, bar);
  };
  if (c == 0)
    for (int s = 0; s < slots && s < n_groups; ++s) issue(s, s);

  const f2 one = bc2(1.0f);
  f2 a1[2], dsc[2], dsh[2];
  {
    uint2 av = make_uint2(0u, 0u);
    if (live) av = __ldg(reinterpret_cast<const uint2*>(scale + (int64_t)b * mod_stride) + c);
    a1[0] = add2(unpack2(av.x), one);
    a1[1] = add2(unpack2(av.y), one);
    dsc[0] = dsc[1] = dsh[0] = dsh[1] = bc2(0.f);
  }
  int slot = 0;
  uint32_t phase = 0;
  for (int g = 0; g < n_groups; ++g) {
    const int rb = r0 + g * kLnRows;
    const uint32_t src = ring0 + slot * slot_bytes + (uint32_t)c * 8u;
    float rs[kLnRows], nmr[kLnRows];
#pragma unroll
    for (int k = 0; k < kLnRows; ++k) {
      const int64_t row = (int64_t)b * N + min(rb + k, r1 - 1);
      rs[k] = rstd_in[row];
      nmr[k] = -mean_in[row] * rs[k];
    }
    mbar_wait(bar0 + slot * 8, phase);
    uint2 xv[kLnRows], gv[kLnRows], rv[kLnRows];
#pragma unroll
    for (int k = 0; k < kLnRows; ++k) {
      const bool ok = live && rb + k < r1;
      xv[k] = ok ? lds64(src + k * row_bytes) : make_uint2(0u, 0u);
      gv[k] = ok ? lds64(src + tensor_bytes + k * row_bytes) : make_uint2(0u, 0u);
      // bf16 zeros without a residual gradient: adding them leaves dx_ln unchanged
      rv[k] = ok && dres != nullptr ? lds64(src + 2u * tensor_bytes + k * row_bytes) : make_uint2(0u, 0u);
    }
    float v[2 * kLnRows];                                   // [k] = sum g, [kLnRows + k] = sum g * xhat
#pragma unroll
    for (int k = 0; k < kLnRows; ++k) {
      const f2 rsk = bc2(rs[k]), nmk = bc2(nmr[k]);
      const f2 g0 = mul2(unpack2(gv[k].x), a1[0]), g1 = mul2(unpack2(gv[k].y), a1[1]);
      const f2 s1 = add2(g0, g1);
      const f2 s2 = fma2(g1, fma2(unpack2(xv[k].y), rsk, nmk), mul2(g0, fma2(unpack2(xv[k].x), rsk, nmk)));
      float lo, hi;
      un2(s1, lo, hi);
      v[k] = lo + hi;
      un2(s2, lo, hi);
      v[kLnRows + k] = lo + hi;
    }
    // warp reduction of 8 values with a halving exchange: 4 + 2 + 1 shuffles leave value (lane >> 2) & 7 ... then 2 more
#pragma unroll
    for (int i = 0; i < 4; ++i) {                           // xor 16: keep 4 of 8
      const bool up = lane & 16;
      const float send = up ? v[i] : v[i + 4], keep = up ? v[i + 4] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {                           // xor 8: keep 2 of 4
      const bool up = lane & 8;
      const float send = up ? v[i] : v[i + 2], keep = up ? v[i + 2] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    {
      const bool up = lane & 4;                             // xor 4: keep 1 of 2
      const float send = up ? v[0] : v[1], keep = up ? v[1] : v[0];
      v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
    // lane L now holds the warp total of value index ((L>>4)&1)*4 + ((L>>3)&1)*2 + ((L>>2)&1)
    const uint32_t red_slot = red0 + (g & 1) * (kLnMaxWarps * 2 * kLnRows * 4);
    if ((lane & 3) == 0) {
      const int vi = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(red_slot + (warp * 2 * kLnRows + vi) * 4), "f"(v[0]) : "memory");
    }
    __syncthreads();
    // every thread has copied this group out of the ring: refill the slot with the group `slots` ahead
    if (c == 0 && g + slots < n_groups) issue(g + slots, slot);
    float mine = 0.f;                                        // lane i < 8: total of value i over the warps, in order
    for (int w = 0; w < nwarps; ++w) mine += lds32f(red_slot + (w * 2 * kLnRows + (lane & 7)) * 4);
#pragma unroll
    for (int k = 0; k < kLnRows; ++k) {
      const float t1 = __shfl_sync(0xffffffffu, mine, k), t2 = __shfl_sync(0xffffffffu, mine, kLnRows + k);
      if (!(live && rb + k < r1)) continue;
      const int64_t row = (int64_t)b * N + rb + k;
      const f2 rsk = bc2(rs[k]), nmk = bc2(nmr[k]);
      const f2 nc1r = bc2(-t1 * inv_c * rs[k]), nc2r = bc2(-t2 * inv_c * rs[k]);
      const uint32_t xw[2] = {xv[k].x, xv[k].y}, gw[2] = {gv[k].x, gv[k].y}, rw[2] = {rv[k].x, rv[k].y};
      uint32_t ow[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const f2 g2 = unpack2(gw[j]);
        const f2 xh = fma2(unpack2(xw[j]), rsk, nmk);
        ow[j] = pack2(add2(fma2(xh, nc2r, fma2(mul2(g2, a1[j]), rsk, nc1r)), unpack2(rw[j])));
        dsc[j] = fma2(g2, xh, dsc[j]);
        dsh[j] = add2(dsh[j], g2);
      }
      asm volatile("st.global.cs.v2.u32 [%0], {%1, %2};" ::"l"(dx + row * C + 4 * c), "r"(ow[0]), "r"(ow[1]) : "memory");
    }
    if (++slot == slots) { slot = 0; phase ^= 1u; }
  }
  if (live) {   // partial of this split: the finish kernel adds gridDim.x partials in index order
    const int64_t o = ((int64_t)split * gridDim.y + b) * C + c * 4;
    float4 q;
    un2(dsc[0], q.x, q.y); un2(dsc[1], q.z, q.w);
    *reinterpret_cast<float4*>(part0 + o) = q;
    un2(dsh[0], q.x, q.y); un2(dsh[1], q.z, q.w);
    *reinterpret_cast<float4*>(part1 + o) = q;
  }
}

// ------------------------------------------------------------------------------------------------
// Column pass: lanes run along C (coalesced 512 B per warp per row), warps stride over the rows of one
// (sample, split); per-thread partial sums -> shared memory over the 8 warps -> partial[split][b][C] fp32.
//   kColLn    sum_n dy * xhat  and  sum_n dy                     (gradients of scale and shift)
//   kColGate  dy = gate * dout written out, sum_n dout * y        (gradient of gate)
// A second tiny kernel adds the splits in fixed order (deterministic) and rounds to bf16.
// ------------------------------------------------------------------------------------------------
enum { kColLn = 0, kColGate = 1 };

template <int MODE>
__global__ void __launch_bounds__(kThreads) column_pass_kernel(
    const uint4* __restrict__ a,       // kColLn: dy      kColGate: dout
    const uint4* __restrict__ bsrc,    // kColLn: x       kColGate: y
    const uint4* __restrict__ gate, int64_t mod_stride_vec, const float* __restrict__ mean_in,
    const float* __restrict__ rstd_in, int N, int nvec, int rows_per_split, uint4* __restrict__ dy_out,
    float* __restrict__ part0, float* __restrict__ part1) {
  __shared__ float red[kWarpsPerBlock][2][32 * 8 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int split = blockIdx.y, b = blockIdx.z;
  const bool active = c < nvec;
  const int r0 = split * rows_per_split;
  const int r1 = min(N, r0 + rows_per_split);
  float acc0[8], acc1[8], gt[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc0[j] = acc1[j] = gt[j] = 0.f;
  if (MODE == kColGate && active) unpack8(ld_cached(gate + (int64_t)b * mod_stride_vec + c), gt);
  if (active) {
    for (int r = r0 + warp; r < r1; r += kWarpsPerBlock) {
      const int64_t row = (int64_t)b * N + r;
      float fa[8], fb[8];
      unpack8(ld_stream(a + row * nvec + c), fa);
      unpack8(ld_stream(bsrc + row * nvec + c), fb);
      if (MODE == kColLn) {
        const float mean = mean_in[row], rstd = rstd_in[row];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc0[j] = fmaf(fa[j], (fb[j] - mean) * rstd, acc0[j]);
          acc1[j] += fa[j];
        }
      } else {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc0[j] = fmaf(fa[j], fb[j], acc0[j]);
          o[j] = gt[j] * fa[j];
        }
        st_stream(dy_out + row * nvec + c, pack8(o));
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    red[warp][0][lane * 8 + j] = acc0[j];
    if (MODE == kColLn) red[warp][1][lane * 8 + j] = acc1[j];
  }
  __syncthreads();
  // 256 threads <-> the block's 256 columns
  const int col = blockIdx.x * 256 + threadIdx.x;
  if (col < nvec * 8) {
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int w = 0; w < kWarpsPerBlock; ++w) {
      s0 += red[w][0][threadIdx.x];
      if (MODE == kColLn) s1 += red[w][1][threadIdx.x];
    }
    const int64_t o = ((int64_t)split * gridDim.z + b) * (nvec * 8) + col;
    part0[o] = s0;
    if (MODE == kColLn) part1[o] = s1;
  }
}

__global__ void __launch_bounds__(256) column_finish_kernel(const float* __restrict__ part0,
                                                            const float* __restrict__ part1, int splits, int64_t bc,
                                                            __nv_bfloat16* __restrict__ out0,
                                                            __nv_bfloat16* __restrict__ out1) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= bc) return;
  float s0 = 0.f, s1 = 0.f;
  for (int s = 0; s < splits; ++s) {
    s0 += part0[(int64_t)s * bc + i];
    if (part1) s1 += part1[(int64_t)s * bc + i];
  }
  out0[i] = __float2bfloat16_rn(s0);
  if (part1) out1[i] = __float2bfloat16_rn(s1);
}

// ------------------------------------------------------------------------------------------------
// colsum: out[c] = sum_r a[r][c] for a [rows, C] bf16 matrix = the bias gradient of a Linear layer (dy.sum(0) in
// autograd: a strided torch reduce at ~2 TB/s).  Same lanes-along-C / warps-over-rows shape as the column pass,
// 4 rows in flight per thread, fp32 partial per (split, column) -> column_finish_kernel (fixed order).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) colsum_kernel(const uint4* __restrict__ a, int64_t rows, int nvec,
                                                          int rows_per_split, float* __restrict__ part) {
  __shared__ float red[kWarpsPerBlock][32 * 8 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_split;
  const int64_t r1 = r0 + rows_per_split < rows ? r0 + rows_per_split : rows;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c < nvec) {
    const uint4* col = a + c;
    int64_t r = r0 + warp;
    for (; r + 3 * kWarpsPerBlock < r1; r += 4 * kWarpsPerBlock) {
      uint4 v[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) v[i] = ld_stream(col + (r + i * kWarpsPerBlock) * nvec);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float f[8];
        unpack8(v[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
    for (; r < r1; r += kWarpsPerBlock) {
      float f[8];
      unpack8(ld_stream(col + r * nvec), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = acc[j];
  __syncthreads();
  const int colx = blockIdx.x * 256 + threadIdx.x;
  if (colx < nvec * 8) {
    float s0 = 0.f;
#pragma unroll
    for (int w = 0; w < kWarpsPerBlock; ++w) s0 += red[w][threadIdx.x];
    part[(int64_t)blockIdx.y * (nvec * 8) + colx] = s0;
  }
}

// finish of colsum: 32 columns x 8 split groups per block (a thread adds every 8th split, the groups are added in
// fixed order through shared memory): 17 dependent loads per thread at 132 splits instead of 132
__global__ void __launch_bounds__(256) colsum_finish_kernel(const float* __restrict__ part, int splits, int C,
                                                            __nv_bfloat16* __restrict__ out) {
  __shared__ float red[8][32];
  const int c = blockIdx.x * 32 + (threadIdx.x & 31), g = threadIdx.x >> 5;
  float s = 0.f;
  if (c < C)
    for (int k = g; k < splits; k += 8) s += part[(int64_t)k * C + c];
  red[g][threadIdx.x & 31] = s;
  __syncthreads();
  if (g == 0 && c < C) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
    out[c] = __float2bfloat16_rn(t);
  }
}

inline int colsum_splits(int64_t rows, int C) {   // ~8 CTAs per SM, at least 32 rows per split
  const int colblocks = (C / 8 + 31) / 32;
  int64_t s = (kNumSMs * 8 + colblocks - 1) / colblocks;
  const int64_t cap = (rows + 31) / 32;
  if (s > cap) s = cap;
  return (int)(s < 1 ? 1 : s);
}

// gate_residual forward: out = res + gate_b * y
__global__ void __launch_bounds__(kThreads) gate_residual_fwd_kernel(
    const uint4* __restrict__ res, const uint4* __restrict__ y, const uint4* __restrict__ gate,
    int64_t mod_stride_vec, int64_t total_vec, int nvec, int64_t vec_per_sample, uint4* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < total_vec; i += (int64_t)gridDim.x * kThreads) {
    const int64_t b = i / vec_per_sample;
    const int c = (int)(i % nvec);
    float r[8], f[8], g[8], o[8];
    unpack8(ld_stream(res + i), r);
    unpack8(ld_stream(y + i), f);
    unpack8(ld_cached(gate + b * mod_stride_vec + c), g);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fmaf(g[j], f[j], r[j]);
    st_stream(out + i, pack8(o));
  }
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

inline int split_count(int N) {  // rows per split >= 64, at most 16 splits
  int s = N / 128;
  return s < 1 ? 1 : (s > 16 ? 16 : s);
}

// ln_modulate backward: upper bound of the row splits per sample (sizes the partial-sum workspace)
inline int ln_bwd_splits(int B, int N, int C) {
  (void)C;
  int s = (kNumSMs * 3) / B;
  const int cap = (N + 3) / 4;
  if (s > cap) s = cap;
  return s < 1 ? 1 : s;
}

// MAXV = 128-bit vectors per lane that cover one row: dispatch on ceil(nvec / 32)
#define YAT_DISPATCH_MAXV(per_lane, ...)                      \
  do {                                                        \
    if (per_lane <= 1)       { constexpr int MAXV = 1;  __VA_ARGS__; } \
    else if (per_lane <= 2)  { constexpr int MAXV = 2;  __VA_ARGS__; } \
    else if (per_lane <= 4)  { constexpr int MAXV = 4;  __VA_ARGS__; } \
    else if (per_lane <= 9)  { constexpr int MAXV = 9;  __VA_ARGS__; } \
    else                     { constexpr int MAXV = 16; __VA_ARGS__; } \
  } while (0)

int check_common(const char* what, int B, int N, int C, int64_t mod_stride) {
  YAT_REQUIRE(B > 0 && N > 0 && C > 0, YAT_ERR_BAD_ARG, "%s: B=%d N=%d C=%d must be positive", what, B, N, C);
  YAT_REQUIRE(C % 8 == 0 && C <= 4096, YAT_ERR_UNSUPPORTED, "%s: C=%d must be a multiple of 8 and <= 4096", what, C);
  YAT_REQUIRE(mod_stride % 8 == 0 && mod_stride >= C, YAT_ERR_BAD_ARG,
              "%s: modulation batch stride %lld must be a multiple of 8 and >= C", what, (long long)mod_stride);
  return YAT_OK;
}

}  // namespace
}  // namespace yat

extern "C" int64_t yat_colsum_workspace_bytes(int64_t rows, int C) {
  return (int64_t)yat::colsum_splits(rows, C) * C * (int64_t)sizeof(float);
}

extern "C" int yat_colsum_bf16(const void* a, int64_t rows, int C, void* out, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(a && out && workspace, YAT_ERR_BAD_ARG, "yat_colsum_bf16: null pointer argument");
  YAT_REQUIRE(rows > 0 && C > 0, YAT_ERR_BAD_ARG, "yat_colsum_bf16: rows=%lld C=%d must be positive", (long long)rows, C);
  YAT_REQUIRE(C % 8 == 0, YAT_ERR_UNSUPPORTED, "yat_colsum_bf16: C=%d must be a multiple of 8", C);
  YAT_REQUIRE(aligned16(a), YAT_ERR_BAD_ARG, "yat_colsum_bf16: the matrix must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int nvec = C / 8, splits = colsum_splits(rows, C);
  const int rows_per_split = (int)((rows + splits - 1) / splits);
  colsum_kernel<<<dim3((nvec + 31) / 32, splits), kThreads, 0, st>>>((const uint4*)a, rows, nvec, rows_per_split,
                                                                    (float*)workspace);
  int rc = check_launch("colsum");
  if (rc != YAT_OK) return rc;
  colsum_finish_kernel<<<(unsigned)((C + 31) / 32), 256, 0, st>>>((const float*)workspace, splits, C, (__nv_bfloat16*)out);
  return check_launch("colsum_finish");
}

// ---- colsum without any allocation: the split partials live in a file-scope device array ------------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).  No counterpart in
// the reference.  2 MiB per GPU that loads the library, held until exit.
// Every partial the finish kernel reads is written by the first kernel of the same call, so stale contents are harmless
// and nothing needs clearing; but ALL calls share the array: two calls that overlap on the GPU (different streams) race.
constexpr int64_t kColsumStaticFloats = 512 * 1024;
static __device__ __align__(256) float g_colsum_scratch[kColsumStaticFloats];

// In host code `g_colsum_scratch` is a host-side stand-in, not a device address: resolve it once per device, with that
// device current, and hand the kernels the resolved pointer.
static int colsum_static_ptr(float** out) {
  constexpr int kMaxDevices = 64;
  static float* ptrs[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  YAT_REQUIRE(e == cudaSuccess && dev >= 0 && dev < kMaxDevices, YAT_ERR_CUDA, "yat_colsum_bf16_static: cudaGetDevice failed (%s)",
              cudaGetErrorString(e));
  float* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    e = cudaGetSymbolAddress((void**)&p, g_colsum_scratch);
    YAT_REQUIRE(e == cudaSuccess && p != nullptr, YAT_ERR_CUDA, "yat_colsum_bf16_static: cudaGetSymbolAddress failed (%s)",
                cudaGetErrorString(e));
    __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);   // racing threads resolve the same address
  }
  *out = p;
  return YAT_OK;
}

extern "C" int64_t yat_colsum_static_max_bytes(void) { return kColsumStaticFloats * (int64_t)sizeof(float); }

extern "C" int yat_colsum_bf16_static(const void* a, int64_t rows, int C, void* out, void* stream) {
  using namespace yat;
  YAT_REQUIRE(a && out, YAT_ERR_BAD_ARG, "yat_colsum_bf16_static: null pointer argument");
  YAT_REQUIRE(rows > 0 && C > 0, YAT_ERR_BAD_ARG, "yat_colsum_bf16_static: rows=%lld C=%d must be positive", (long long)rows, C);
  YAT_REQUIRE(C % 8 == 0, YAT_ERR_UNSUPPORTED, "yat_colsum_bf16_static: C=%d must be a multiple of 8", C);
  YAT_REQUIRE(aligned16(a), YAT_ERR_BAD_ARG, "yat_colsum_bf16_static: the matrix must be 16-byte aligned");
  const int nvec = C / 8, splits = colsum_splits(rows, C);
  YAT_REQUIRE((int64_t)splits * C <= kColsumStaticFloats, YAT_ERR_UNSUPPORTED,
              "yat_colsum_bf16_static: rows=%lld C=%d needs %lld B of scratch, the built-in array holds %lld B; use "
              "yat_colsum_bf16 with a caller-owned workspace", (long long)rows, C,
              (long long)((int64_t)splits * C * sizeof(float)), (long long)(kColsumStaticFloats * sizeof(float)));
  float* scratch = nullptr;
  int rc = colsum_static_ptr(&scratch);
  if (rc != YAT_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int rows_per_split = (int)((rows + splits - 1) / splits);
  colsum_kernel<<<dim3((nvec + 31) / 32, splits), kThreads, 0, st>>>((const uint4*)a, rows, nvec, rows_per_split, scratch);
  rc = check_launch("colsum (static scratch)");
  if (rc != YAT_OK) return rc;
  colsum_finish_kernel<<<(unsigned)((C + 31) / 32), 256, 0, st>>>(scratch, splits, C, (__nv_bfloat16*)out);
  return check_launch("colsum_finish (static scratch)");
}

extern "C" int64_t yat_block_workspace_bytes(int B, int N, int C) {
  const int a = yat::split_count(N), b = yat::ln_bwd_splits(B, N, C);
  return (int64_t)2 * (a > b ? a : b) * B * C * (int64_t)sizeof(float);
}

extern "C" int yat_ln_modulate_fwd(const void* x, const void* scale, const void* shift, int64_t mod_batch_stride,
                                   int B, int N, int C, float eps, void* y, float* mean, float* rstd, void* stream) {
  return yat_add_ln_modulate_fwd(x, nullptr, scale, shift, mod_batch_stride, B, N, C, eps, nullptr, y, mean, rstd, stream);
}

extern "C" int yat_add_ln_modulate_fwd(const void* x, const void* x2, const void* scale, const void* shift,
                                       int64_t mod_batch_stride, int B, int N, int C, float eps, void* xsum, void* y,
                                       float* mean, float* rstd, void* stream) {
  using namespace yat;
  YAT_REQUIRE(x && scale && shift && y, YAT_ERR_BAD_ARG, "yat_ln_modulate_fwd: null pointer argument");
  YAT_REQUIRE((x2 == nullptr) == (xsum == nullptr), YAT_ERR_BAD_ARG, "yat_add_ln_modulate_fwd: x2 and xsum go together");
  YAT_REQUIRE(!x2 || (aligned16(x2) && aligned16(xsum)), YAT_ERR_BAD_ARG,
              "yat_add_ln_modulate_fwd: tensors must be 16-byte aligned");
  YAT_REQUIRE((mean == nullptr) == (rstd == nullptr), YAT_ERR_BAD_ARG, "yat_ln_modulate_fwd: mean and rstd go together");
  int rc = check_common("yat_ln_modulate_fwd", B, N, C, mod_batch_stride);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(x) && aligned16(scale) && aligned16(shift) && aligned16(y), YAT_ERR_BAD_ARG,
              "yat_ln_modulate_fwd: tensors must be 16-byte aligned");
  const int nvec = C / 8, rows = B * N;
  const int grid = (rows + kWarpsPerBlock - 1) / kWarpsPerBlock;
  if (x2)
    YAT_DISPATCH_MAXV((nvec + 31) / 32, ln_modulate_fwd_kernel<MAXV, true><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        (const uint4*)x, (const uint4*)x2, (const uint4*)scale, (const uint4*)shift, mod_batch_stride / 8, rows, N, nvec,
        1.0f / C, eps, (uint4*)xsum, (uint4*)y, mean, rstd));
  else
    YAT_DISPATCH_MAXV((nvec + 31) / 32, ln_modulate_fwd_kernel<MAXV, false><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        (const uint4*)x, nullptr, (const uint4*)scale, (const uint4*)shift, mod_batch_stride / 8, rows, N, nvec,
        1.0f / C, eps, nullptr, (uint4*)y, mean, rstd));
  return check_launch("ln_modulate_fwd");
}

extern "C" int yat_ln_modulate_bwd(const void* x, const void* dy, const void* scale, int64_t mod_batch_stride,
                                   const float* mean, const float* rstd, int B, int N, int C, void* dx, void* dscale,
                                   void* dshift, void* workspace, void* stream) {
  return yat_ln_modulate_bwd_res(x, dy, nullptr, scale, mod_batch_stride, mean, rstd, B, N, C, dx, dscale, dshift,
                                 workspace, stream);
}

extern "C" int yat_ln_modulate_bwd_res(const void* x, const void* dy, const void* dres, const void* scale,
                                       int64_t mod_batch_stride, const float* mean, const float* rstd, int B, int N,
                                       int C, void* dx, void* dscale, void* dshift, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(!dres || aligned16(dres), YAT_ERR_BAD_ARG, "yat_ln_modulate_bwd_res: dres must be 16-byte aligned");
  YAT_REQUIRE(x && dy && scale && mean && rstd && dx && dscale && dshift && workspace, YAT_ERR_BAD_ARG,
              "yat_ln_modulate_bwd: null pointer argument");
  int rc = check_common("yat_ln_modulate_bwd", B, N, C, mod_batch_stride);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(x) && aligned16(dy) && aligned16(scale) && aligned16(dx), YAT_ERR_BAD_ARG,
              "yat_ln_modulate_bwd: tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int nvec = C / 8, rows = B * N;
  (void)rows;
  const int splits = ln_bwd_splits(B, N, C);
  float* part0 = (float*)workspace;
  {
    const int threads = (C / 4 + 31) / 32 * 32;
    const int slots = ln_bwd_slots(C);
    YAT_REQUIRE(slots >= 2 && threads <= 1024, YAT_ERR_UNSUPPORTED,
                "yat_ln_modulate_bwd: C=%d does not fit the shared-memory row ring", C);
    const int64_t smem = ln_bwd_smem_bytes(slots, C);
    int sp = kNumSMs / B;                                          // one CTA per SM
    const int cap = (N + kLnRows - 1) / kLnRows;
    sp = sp > cap ? cap : sp;
    sp = sp > splits ? splits : sp;                                // the workspace is sized for `splits` partials
    sp = sp < 1 ? 1 : sp;
    const int rps = (N + sp - 1) / sp;
    float* p1 = part0 + (int64_t)sp * B * C;
    static bool attr_set = false;
    if (!attr_set) {
      cudaError_t e =
          cudaFuncSetAttribute(ln_modulate_bwd_ring_kernel<640>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
      if (e == cudaSuccess)
        e = cudaFuncSetAttribute(ln_modulate_bwd_ring_kernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
      YAT_REQUIRE(e == cudaSuccess, YAT_ERR_CUDA, "yat_ln_modulate_bwd: shared memory opt-in failed (%s)",
                  cudaGetErrorString(e));
      attr_set = true;
    }
    if (threads <= 640)    // SANA-1.6B: C = 2240 -> 560 columns-of-4, 576 threads, no register spills
      ln_modulate_bwd_ring_kernel<640><<<dim3(sp, B), threads, smem, st>>>(
          (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, (const __nv_bfloat16*)scale, mod_batch_stride, mean, rstd, N,
          C, slots, rps, 1.0f / C, (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, part0, p1);
    else
      ln_modulate_bwd_ring_kernel<1024><<<dim3(sp, B), threads, smem, st>>>(
          (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, (const __nv_bfloat16*)scale, mod_batch_stride, mean, rstd, N,
          C, slots, rps, 1.0f / C, (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, part0, p1);
    rc = check_launch("ln_modulate_bwd_cols");
    if (rc != YAT_OK) return rc;
    const int64_t bc2_ = (int64_t)B * C;
    column_finish_kernel<<<(unsigned)((bc2_ + 255) / 256), 256, 0, st>>>(part0, p1, sp, bc2_, (__nv_bfloat16*)dscale,
                                                                        (__nv_bfloat16*)dshift);
    return check_launch("ln_modulate_bwd_finish");
  }
}

extern "C" int yat_gate_residual_fwd(const void* res, const void* y, const void* gate, int64_t mod_batch_stride, int B,
                                     int N, int C, void* out, void* stream) {
  using namespace yat;
  YAT_REQUIRE(res && y && gate && out, YAT_ERR_BAD_ARG, "yat_gate_residual_fwd: null pointer argument");
  int rc = check_common("yat_gate_residual_fwd", B, N, C, mod_batch_stride);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(res) && aligned16(y) && aligned16(gate) && aligned16(out), YAT_ERR_BAD_ARG,
              "yat_gate_residual_fwd: tensors must be 16-byte aligned");
  const int nvec = C / 8;
  const int64_t total = (int64_t)B * N * nvec;
  const int64_t want = (total + kThreads - 1) / kThreads;
  const int grid = (int)(want < kNumSMs * 16 ? want : kNumSMs * 16);
  gate_residual_fwd_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(
      (const uint4*)res, (const uint4*)y, (const uint4*)gate, mod_batch_stride / 8, total, nvec, (int64_t)N * nvec,
      (uint4*)out);
  return check_launch("gate_residual_fwd");
}

extern "C" int yat_gate_residual_bwd(const void* dout, const void* y, const void* gate, int64_t mod_batch_stride, int B,
                                     int N, int C, void* dy, void* dgate, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(dout && y && gate && dy && dgate && workspace, YAT_ERR_BAD_ARG,
              "yat_gate_residual_bwd: null pointer argument");
  int rc = check_common("yat_gate_residual_bwd", B, N, C, mod_batch_stride);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(dout) && aligned16(y) && aligned16(gate) && aligned16(dy), YAT_ERR_BAD_ARG,
              "yat_gate_residual_bwd: tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int nvec = C / 8;
  const int splits = split_count(N), rows_per_split = (N + splits - 1) / splits;
  float* part0 = (float*)workspace;
  const dim3 g((nvec + 31) / 32, splits, B);
  column_pass_kernel<kColGate><<<g, kThreads, 0, st>>>((const uint4*)dout, (const uint4*)y, (const uint4*)gate,
                                                       mod_batch_stride / 8, nullptr, nullptr, N, nvec, rows_per_split,
                                                       (uint4*)dy, part0, nullptr);
  const int64_t bc = (int64_t)B * C;
  column_finish_kernel<<<(unsigned)((bc + 255) / 256), 256, 0, st>>>(part0, nullptr, splits, bc, (__nv_bfloat16*)dgate,
                                                                    nullptr);
  return check_launch("gate_residual_bwd");
}
