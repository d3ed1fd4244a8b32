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

namespace yat {
namespace {

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
template <int MAXV>
__global__ void __launch_bounds__(kThreads) ln_modulate_fwd_kernel(
    const uint4* __restrict__ x, const uint4* __restrict__ scale, const uint4* __restrict__ shift,
    int64_t mod_stride_vec, int rows, int N, int nvec, float inv_c, float eps, uint4* __restrict__ y,
    float* __restrict__ mean_out, float* __restrict__ rstd_out) {
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

// ln_modulate backward, row part: dx = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * (1 + scale)
template <int MAXV>
__global__ void __launch_bounds__(kThreads) ln_modulate_bwd_dx_kernel(
    const uint4* __restrict__ x, const uint4* __restrict__ dy, const uint4* __restrict__ scale,
    int64_t mod_stride_vec, const float* __restrict__ mean_in, const float* __restrict__ rstd_in, int rows, int N,
    int nvec, float inv_c, uint4* __restrict__ dx) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int b = row / N;
  const uint4* xr = x + (int64_t)row * nvec;
  const uint4* gr = dy + (int64_t)row * nvec;
  const uint4* sc = scale + (int64_t)b * mod_stride_vec;
  const float mean = mean_in[row], rstd = rstd_in[row];
  uint4 xv[MAXV], gv[MAXV];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      xv[i] = ld_stream(xr + c);
      gv[i] = ld_stream(gr + c);
    }
  }
  // pass 1: g (kept packed as bf16x2-free fp32 is too register hungry: recompute in pass 2), the two row means
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      float f[8], g[8], a[8];
      unpack8(xv[i], f);
      unpack8(gv[i], g);
      unpack8(ld_cached(sc + c), a);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float gg = g[j] * (1.0f + a[j]);
        s1 += gg;
        s2 = fmaf(gg, (f[j] - mean) * rstd, s2);
      }
    }
  }
  const float c1 = warp_sum(s1) * inv_c, c2 = warp_sum(s2) * inv_c;
  uint4* dr = dx + (int64_t)row * nvec;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      float f[8], g[8], a[8], o[8];
      unpack8(xv[i], f);
      unpack8(gv[i], g);
      unpack8(ld_cached(sc + c), a);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xh = (f[j] - mean) * rstd;
        o[j] = rstd * (g[j] * (1.0f + a[j]) - c1 - xh * c2);
      }
      st_stream(dr + c, pack8(o));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// ln_modulate backward, ONE pass: dx (row part, as above) and the per-sample column sums
//   dscale[b][c] = sum_n dy * xhat      dshift[b][c] = sum_n dy
// from the same registers, so x and dy are read once (6·C B per token instead of 10·C).  CTA = (row split, sample),
// 8 warps = 4 warp pairs, one row per pair at a time: the two warps of a pair own alternate 32-vector blocks of the
// row (x and dy of their half live in registers, 2 * MAXH 128-bit loads in flight per lane) and exchange their two
// partial row sums through shared memory with a 64-thread named barrier (slots alternate by row, so one barrier per
// row).  Every pair owns a private fp32 accumulator image of the two [C] rows in shared memory, laid out as planes of
// 16 bytes per 8-column vector so that lane-contiguous accesses are conflict-free, and touched only by the lane that
// owns the vector.  Arithmetic on packed fp32 pairs (FFMA2).  Pairs are added in fixed order, splits by
// column_finish_kernel -> bit-identical run to run.
// ------------------------------------------------------------------------------------------------
constexpr int kLnBwdPairs = 4;
constexpr int kLnBwdThreads = kLnBwdPairs * 64;

template <int MAXH>
__global__ void __launch_bounds__(kLnBwdThreads, MAXH <= 5 ? 3 : 1) ln_modulate_bwd_fused_kernel(
    const uint4* __restrict__ x, const uint4* __restrict__ dy, const uint4* __restrict__ scale,
    int64_t mod_stride_vec, const float* __restrict__ mean_in, const float* __restrict__ rstd_in, int N, int nvec,
    int rows_per_split, float inv_c, uint4* __restrict__ dx, float* __restrict__ part0, float* __restrict__ part1) {
  extern __shared__ __align__(16) ulonglong2 acc_smem[];          // [pair][kind: dscale, dshift][plane][nvec] x 4 floats
  __shared__ float xchg[kLnBwdPairs][2][2][2];                    // [pair][row parity][half][s1, s2]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int pair = warp >> 1, half = warp & 1;
  const int split = blockIdx.x, b = blockIdx.y;
  ulonglong2* acc = acc_smem + (size_t)pair * 4 * nvec;
  for (int i = (half * 32 + lane); i < 4 * nvec; i += 64) acc[i] = make_ulonglong2(0ull, 0ull);   // fp32 +0.0 pairs
  asm volatile("bar.sync %0, 64;" ::"r"(pair + 1) : "memory");
  const int r0 = split * rows_per_split;
  const int r1 = min(N, r0 + rows_per_split);
  const uint4* sc = scale + (int64_t)b * mod_stride_vec;
  const f2 one = bc2(1.0f);
  int parity = 0;
  for (int r = r0 + pair; r < r1; r += kLnBwdPairs, parity ^= 1) {
    const int64_t row = (int64_t)b * N + r;
    const uint4* xr = x + row * nvec;
    const uint4* gr = dy + row * nvec;
    if (lane == 0 && r + 2 * kLnBwdPairs < r1) {                   // pull the pair's row after next into L2: no registers held
      const uint4* nxt = (half ? gr : xr) + (int64_t)2 * kLnBwdPairs * nvec;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(nxt), "r"(nvec * 16) : "memory");
    }
    uint4 xv[MAXH], gv[MAXH];
#pragma unroll
    for (int i = 0; i < MAXH; ++i) {
      const int c = (2 * i + half) * 32 + lane;
      if (c < nvec) {
        xv[i] = ld_stream(xr + c);
        gv[i] = ld_stream(gr + c);
      }
    }
    const float mean = mean_in[row], rstd = rstd_in[row];
    const f2 rs = bc2(rstd), nmr = bc2(-mean * rstd);              // xhat = x * rstd - mean * rstd
    f2 s1 = bc2(0.f), s2 = bc2(0.f);
#pragma unroll
    for (int i = 0; i < MAXH; ++i) {
      const int c = (2 * i + half) * 32 + lane;
      if (c < nvec) {
        const uint4 av = ld_cached(sc + c);
        const uint32_t xw[4] = {xv[i].x, xv[i].y, xv[i].z, xv[i].w}, gw[4] = {gv[i].x, gv[i].y, gv[i].z, gv[i].w},
                       aw[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const f2 gg = mul2(unpack2(gw[j]), add2(unpack2(aw[j]), one));
          s1 = add2(s1, gg);
          s2 = fma2(gg, fma2(unpack2(xw[j]), rs, nmr), s2);
        }
      }
    }
    float s1a, s1b, s2a, s2b;
    un2(s1, s1a, s1b);
    un2(s2, s2a, s2b);
    const float h1 = warp_sum(s1a + s1b), h2 = warp_sum(s2a + s2b);
    if (lane == 0) {
      xchg[pair][parity][half][0] = h1;
      xchg[pair][parity][half][1] = h2;
    }
    asm volatile("bar.sync %0, 64;" ::"r"(pair + 1) : "memory");
    const float c1 = (xchg[pair][parity][0][0] + xchg[pair][parity][1][0]) * inv_c;
    const float c2 = (xchg[pair][parity][0][1] + xchg[pair][parity][1][1]) * inv_c;
    const f2 nc1r = bc2(-c1 * rstd), nc2r = bc2(-c2 * rstd);       // dx = gg * rstd - c1 * rstd - xhat * c2 * rstd
    uint4* dr = dx + row * nvec;
#pragma unroll
    for (int i = 0; i < MAXH; ++i) {
      const int c = (2 * i + half) * 32 + lane;
      if (c < nvec) {
        const uint4 av = ld_cached(sc + c);
        const uint32_t xw[4] = {xv[i].x, xv[i].y, xv[i].z, xv[i].w}, gw[4] = {gv[i].x, gv[i].y, gv[i].z, gv[i].w},
                       aw[4] = {av.x, av.y, av.z, av.w};
        uint32_t ow[4];
        f2 dsc[4], dsh[4];
        {
          const ulonglong2 t0 = acc[0 * nvec + c], t1 = acc[1 * nvec + c], t2 = acc[2 * nvec + c], t3 = acc[3 * nvec + c];
          dsc[0].v = t0.x; dsc[1].v = t0.y; dsc[2].v = t1.x; dsc[3].v = t1.y;
          dsh[0].v = t2.x; dsh[1].v = t2.y; dsh[2].v = t3.x; dsh[3].v = t3.y;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const f2 g = unpack2(gw[j]);
          const f2 xh = fma2(unpack2(xw[j]), rs, nmr);
          const f2 gg = mul2(g, add2(unpack2(aw[j]), one));
          ow[j] = pack2(fma2(xh, nc2r, fma2(gg, rs, nc1r)));
          dsc[j] = fma2(g, xh, dsc[j]);
          dsh[j] = add2(dsh[j], g);
        }
        st_stream(dr + c, make_uint4(ow[0], ow[1], ow[2], ow[3]));
        acc[0 * nvec + c] = make_ulonglong2(dsc[0].v, dsc[1].v);
        acc[1 * nvec + c] = make_ulonglong2(dsc[2].v, dsc[3].v);
        acc[2 * nvec + c] = make_ulonglong2(dsh[0].v, dsh[1].v);
        acc[3 * nvec + c] = make_ulonglong2(dsh[2].v, dsh[3].v);
      }
    }
  }
  __syncthreads();
  const float* af = reinterpret_cast<const float*>(acc_smem);
  const int C = nvec * 8;
  for (int col = threadIdx.x; col < C; col += kLnBwdThreads) {
    const int c = col >> 3, plane = (col >> 2) & 1, j = col & 3;
    float t0 = 0.f, t1 = 0.f;
#pragma unroll
    for (int w = 0; w < kLnBwdPairs; ++w) {
      t0 += af[(((size_t)w * 4 + plane) * nvec + c) * 4 + j];
      t1 += af[(((size_t)w * 4 + 2 + plane) * nvec + c) * 4 + j];
    }
    const int64_t o = ((int64_t)split * gridDim.y + b) * C + col;
    part0[o] = t0;
    part1[o] = t1;
  }
}

// ------------------------------------------------------------------------------------------------
// ln_modulate backward, column-ownership form (the one launched): a thread owns ONE 8-column vector for the whole CTA
// = (row split, sample), so (1 + scale) is unpacked once into registers and the dscale / dshift column sums are plain
// register accumulators written out at the end (no shared-memory image, no reduction over warps).  Rows go four at a
// time: 8 x 128-bit loads in flight per thread, the 8 partial row sums (4 rows x {sum g, sum g*xhat}) are reduced in
// the warp with a halving exchange (9 shuffles), across the warps through shared memory (every thread adds the warps
// in the same order -> identical statistics everywhere, deterministic), ONE __syncthreads per 4 rows (the slot
// alternates).  x and dy are read once; fp32 pairs (FFMA2).
// ------------------------------------------------------------------------------------------------
constexpr int kLnRows = 4;
constexpr int kLnMaxWarps = 16;   // C <= 4096 -> nvec <= 512 threads

template <int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB) ln_modulate_bwd_cols_kernel(
    const uint4* __restrict__ x, const uint4* __restrict__ dy, const uint4* __restrict__ scale,
    int64_t mod_stride_vec, const float* __restrict__ mean_in, const float* __restrict__ rstd_in, int N, int nvec,
    int rows_per_split, float inv_c, uint4* __restrict__ dx, float* __restrict__ part0, float* __restrict__ part1) {
  __shared__ float red[2][kLnMaxWarps][2 * kLnRows];
  const int c = threadIdx.x, lane = c & 31, warp = c >> 5, nwarps = (blockDim.x + 31) >> 5;
  const int split = blockIdx.x, b = blockIdx.y;
  const bool live = c < nvec;
  const int r0 = split * rows_per_split;
  const int r1 = min(N, r0 + rows_per_split);
  const f2 one = bc2(1.0f);
  f2 a1[4], dsc[4], dsh[4];
  {
    const uint4 av = live ? ld_cached(scale + (int64_t)b * mod_stride_vec + c) : make_uint4(0u, 0u, 0u, 0u);
    const uint32_t aw[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      a1[j] = add2(unpack2(aw[j]), one);
      dsc[j] = bc2(0.f);
      dsh[j] = bc2(0.f);
    }
  }
  int slot = 0;
  for (int rb = r0; rb < r1; rb += kLnRows, slot ^= 1) {
    uint4 xv[kLnRows], gv[kLnRows];
    float rs[kLnRows], nmr[kLnRows];
#pragma unroll
    for (int k = 0; k < kLnRows; ++k) {
      const bool ok = live && rb + k < r1;
      const int64_t row = (int64_t)b * N + min(rb + k, r1 - 1);
      xv[k] = ok ? ld_stream(x + row * nvec + c) : make_uint4(0u, 0u, 0u, 0u);
      gv[k] = ok ? ld_stream(dy + row * nvec + c) : make_uint4(0u, 0u, 0u, 0u);
      rs[k] = rstd_in[row];
      nmr[k] = -mean_in[row] * rs[k];
    }
    float v[2 * kLnRows];                                   // [k] = sum g, [kLnRows + k] = sum g * xhat
#pragma unroll
    for (int k = 0; k < kLnRows; ++k) {
      const uint32_t xw[4] = {xv[k].x, xv[k].y, xv[k].z, xv[k].w}, gw[4] = {gv[k].x, gv[k].y, gv[k].z, gv[k].w};
      const f2 rsk = bc2(rs[k]), nmk = bc2(nmr[k]);
      f2 s1 = bc2(0.f), s2 = bc2(0.f);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const f2 gg = mul2(unpack2(gw[j]), a1[j]);
        s1 = add2(s1, gg);
        s2 = fma2(gg, fma2(unpack2(xw[j]), rsk, nmk), s2);
      }
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
    if ((lane & 3) == 0) red[slot][warp][((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1)] = v[0];
    __syncthreads();
    float tot[2 * kLnRows];
#pragma unroll
    for (int i = 0; i < 2 * kLnRows; ++i) tot[i] = 0.f;
    for (int w = 0; w < nwarps; ++w) {
      const float4 p0 = *reinterpret_cast<const float4*>(&red[slot][w][0]);
      const float4 p1 = *reinterpret_cast<const float4*>(&red[slot][w][4]);
      tot[0] += p0.x; tot[1] += p0.y; tot[2] += p0.z; tot[3] += p0.w;
      tot[4] += p1.x; tot[5] += p1.y; tot[6] += p1.z; tot[7] += p1.w;
    }
#pragma unroll
    for (int k = 0; k < kLnRows; ++k) {
      if (!(live && rb + k < r1)) continue;
      const int64_t row = (int64_t)b * N + rb + k;
      const uint32_t xw[4] = {xv[k].x, xv[k].y, xv[k].z, xv[k].w}, gw[4] = {gv[k].x, gv[k].y, gv[k].z, gv[k].w};
      const f2 rsk = bc2(rs[k]), nmk = bc2(nmr[k]);
      const f2 nc1r = bc2(-tot[k] * inv_c * rs[k]), nc2r = bc2(-tot[kLnRows + k] * inv_c * rs[k]);
      uint32_t ow[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const f2 g = unpack2(gw[j]);
        const f2 xh = fma2(unpack2(xw[j]), rsk, nmk);
        ow[j] = pack2(fma2(xh, nc2r, fma2(mul2(g, a1[j]), rsk, nc1r)));
        dsc[j] = fma2(g, xh, dsc[j]);
        dsh[j] = add2(dsh[j], g);
      }
      st_stream(dx + row * nvec + c, make_uint4(ow[0], ow[1], ow[2], ow[3]));
    }
  }
  if (live) {
    const int64_t o = ((int64_t)split * gridDim.y + b) * (nvec * 8) + c * 8;
    float4 q0, q1;
    un2(dsc[0], q0.x, q0.y); un2(dsc[1], q0.z, q0.w); un2(dsc[2], q1.x, q1.y); un2(dsc[3], q1.z, q1.w);
    *reinterpret_cast<float4*>(part0 + o) = q0;
    *reinterpret_cast<float4*>(part0 + o + 4) = q1;
    un2(dsh[0], q0.x, q0.y); un2(dsh[1], q0.z, q0.w); un2(dsh[2], q1.x, q1.y); un2(dsh[3], q1.z, q1.w);
    *reinterpret_cast<float4*>(part1 + o) = q0;
    *reinterpret_cast<float4*>(part1 + o + 4) = q1;
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

// fused ln_modulate backward: one wave of CTAs (B * splits <= SMs * resident CTAs), at least one row per warp pair
inline int ln_bwd_smem(int C) { return kLnBwdPairs * 2 * C * (int)sizeof(float); }
inline int ln_bwd_splits(int B, int N, int C) {
  int per_sm = (227 * 1024) / (ln_bwd_smem(C) + 1024);
  per_sm = per_sm < 1 ? 1 : (per_sm > 3 ? 3 : per_sm);
  int s = (kNumSMs * per_sm) / B;
  const int cap = (N + kLnBwdPairs - 1) / kLnBwdPairs;
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

extern "C" int64_t yat_block_workspace_bytes(int B, int N, int C) {
  const int a = yat::split_count(N), b = yat::ln_bwd_splits(B, N, C);
  return (int64_t)2 * (a > b ? a : b) * B * C * (int64_t)sizeof(float);
}

extern "C" int yat_ln_modulate_fwd(const void* x, const void* scale, const void* shift, int64_t mod_batch_stride,
                                   int B, int N, int C, float eps, void* y, float* mean, float* rstd, void* stream) {
  using namespace yat;
  YAT_REQUIRE(x && scale && shift && y, YAT_ERR_BAD_ARG, "yat_ln_modulate_fwd: null pointer argument");
  YAT_REQUIRE((mean == nullptr) == (rstd == nullptr), YAT_ERR_BAD_ARG, "yat_ln_modulate_fwd: mean and rstd go together");
  int rc = check_common("yat_ln_modulate_fwd", B, N, C, mod_batch_stride);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(x) && aligned16(scale) && aligned16(shift) && aligned16(y), YAT_ERR_BAD_ARG,
              "yat_ln_modulate_fwd: tensors must be 16-byte aligned");
  const int nvec = C / 8, rows = B * N;
  const int grid = (rows + kWarpsPerBlock - 1) / kWarpsPerBlock;
  YAT_DISPATCH_MAXV((nvec + 31) / 32, ln_modulate_fwd_kernel<MAXV><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
      (const uint4*)x, (const uint4*)scale, (const uint4*)shift, mod_batch_stride / 8, rows, N, nvec, 1.0f / C, eps,
      (uint4*)y, mean, rstd));
  return check_launch("ln_modulate_fwd");
}

extern "C" int yat_ln_modulate_bwd(const void* x, const void* dy, const void* scale, int64_t mod_batch_stride,
                                   const float* mean, const float* rstd, int B, int N, int C, void* dx, void* dscale,
                                   void* dshift, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(x && dy && scale && mean && rstd && dx && dscale && dshift && workspace, YAT_ERR_BAD_ARG,
              "yat_ln_modulate_bwd: null pointer argument");
  int rc = check_common("yat_ln_modulate_bwd", B, N, C, mod_batch_stride);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(x) && aligned16(dy) && aligned16(scale) && aligned16(dx), YAT_ERR_BAD_ARG,
              "yat_ln_modulate_bwd: tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int nvec = C / 8, rows = B * N;
  (void)rows;
  const int splits = ln_bwd_splits(B, N, C), rows_per_split = (N + splits - 1) / splits;
  const int smem = ln_bwd_smem(C);
  float* part0 = (float*)workspace;
  float* part1 = part0 + (int64_t)splits * B * C;
#define YAT_LN_BWD_LAUNCH(MAXH)                                                                                       \
  do {                                                                                                               \
    static bool attr_set = false;                                                                                    \
    if (!attr_set) {                                                                                                 \
      cudaFuncSetAttribute(ln_modulate_bwd_fused_kernel<MAXH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 144 * 1024); \
      attr_set = true;                                                                                               \
    }                                                                                                                \
    ln_modulate_bwd_fused_kernel<MAXH><<<dim3(splits, B), kLnBwdThreads, smem, st>>>(                                 \
        (const uint4*)x, (const uint4*)dy, (const uint4*)scale, mod_batch_stride / 8, mean, rstd, N, nvec,           \
        rows_per_split, 1.0f / C, (uint4*)dx, part0, part1);                                                         \
  } while (0)
  static const bool use_pairs = getenv("YAT_LN_BWD_PAIRS") != nullptr;   // development switch: the warp-pair kernel
  if (!use_pairs) {
    const int threads = (nvec + 31) / 32 * 32;
    const int per_sm = threads <= 320 ? 2 : 1;                     // resident CTAs (launch bounds of the two instantiations)
    int sp = (kNumSMs * per_sm) / B;
    const int cap = (N + kLnRows - 1) / kLnRows;
    sp = sp > cap ? cap : sp;
    sp = sp > splits ? splits : sp;                                // the workspace is sized for `splits`
    sp = sp < 1 ? 1 : sp;
    const int rps = (N + sp - 1) / sp;
    float* p1 = part0 + (int64_t)sp * B * C;
    if (threads <= 320)
      ln_modulate_bwd_cols_kernel<320, 2><<<dim3(sp, B), threads, 0, st>>>(
          (const uint4*)x, (const uint4*)dy, (const uint4*)scale, mod_batch_stride / 8, mean, rstd, N, nvec, rps, 1.0f / C,
          (uint4*)dx, part0, p1);
    else
      ln_modulate_bwd_cols_kernel<512, 1><<<dim3(sp, B), threads, 0, st>>>(
          (const uint4*)x, (const uint4*)dy, (const uint4*)scale, mod_batch_stride / 8, mean, rstd, N, nvec, rps, 1.0f / C,
          (uint4*)dx, part0, p1);
    rc = check_launch("ln_modulate_bwd_cols");
    if (rc != YAT_OK) return rc;
    const int64_t bc2_ = (int64_t)B * C;
    column_finish_kernel<<<(unsigned)((bc2_ + 255) / 256), 256, 0, st>>>(part0, p1, sp, bc2_, (__nv_bfloat16*)dscale,
                                                                        (__nv_bfloat16*)dshift);
    return check_launch("ln_modulate_bwd_finish");
  }
  const int per_half = ((nvec + 31) / 32 + 1) / 2;                 // 32-vector blocks per warp of a pair
  if (per_half <= 1)      YAT_LN_BWD_LAUNCH(1);
  else if (per_half <= 2) YAT_LN_BWD_LAUNCH(2);
  else if (per_half <= 5) YAT_LN_BWD_LAUNCH(5);
  else                    YAT_LN_BWD_LAUNCH(8);
#undef YAT_LN_BWD_LAUNCH
  rc = check_launch("ln_modulate_bwd_fused");
  if (rc != YAT_OK) return rc;
  const int64_t bc = (int64_t)B * C;
  column_finish_kernel<<<(unsigned)((bc + 255) / 256), 256, 0, st>>>(part0, part1, splits, bc, (__nv_bfloat16*)dscale,
                                                                    (__nv_bfloat16*)dshift);
  return check_launch("ln_modulate_bwd_finish");
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
