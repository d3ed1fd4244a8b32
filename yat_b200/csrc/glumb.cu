// Fused middle of SANA's GLUMBConv feed-forward (the reference runs it as
// nn.SiLU + cuDNN depthwise conv + bias + chunk + SiLU + mul, utils/patched_sana_transformer.py GLUMBConv
// via utils/patch_sana_attention_layers.py:106-111):
//
//   u   [B, H, W, 2*Hd] bf16   output of conv_inverted (1x1, run as a GEMM on token-major data)
//   s   = silu(u)                                   (kept in fp32; the reference rounds it to bf16)
//   c   = depthwise3x3(s; w[2*Hd][3][3]) + b        (zero padding 1, fp32 accumulation)
//   mid = c[:Hd] * silu(c[Hd:])        [B, H, W, Hd] bf16   -> input of conv_point (1x1 GEMM)
//
// Token-major (= NHWC) throughout: channels are the contiguous axis, so every global access is a
// full 128-byte line per 64 channels and there are no NCHW<->NHWC permutes around the convolutions.
//
// One CTA = (sample b, 8 image columns, 64 channel pairs) sweeping the rows y = 0..H-1.  A 4-row ring in
// shared memory holds silu(u) in fp32 (each u element is loaded and activated once per tile, one halo
// column each side); warp w owns column w, lane l the channel pairs (c0+2l, c0+2l+1) and their two gate
// channels.  Each thread keeps its 3x3x4 window and its 9x4 weights in registers and slides down the
// column: 6 shared-memory loads and 36 FMAs per row, no cross-lane traffic.  The global loads of row y+2
// are issued before the stencil of row y and committed after it.  The kernels are instruction-bound
// (~40 thread instructions per output), which is why the activation costs a single MUFU op.
//
// Backward (two kernels, same sweep):
//   dc kernel   recompute c; dxa = dmid * silu(g), dg = dmid * xa * silu'(g)          -> dc [B,H,W,2Hd]
//               dw[k] += s[window tap k] * dc, db += dc; per-CTA partial sums -> fixed-order finish
//   du kernel   ds = depthwise3x3^T(dc), du = ds * silu'(u)
#include "common.cuh"

namespace yat {
namespace {

constexpr int kThreads = 256;
constexpr int kRing = 4;
constexpr int kTW = 8;                                       // image columns per CTA (= warps)
constexpr int kRowFloats = (kTW + 2) * 128;                  // one ring row: [kTW + 2][128] fp32

// silu through ONE special-function op:  sigmoid(x) = 0.5 + 0.5 tanh(x / 2)  =>  silu(x) = h + h tanh(h), h = x / 2.
// tanh.approx.f32 has 2^-11 relative error, i.e. |error(silu)| <= 2.5e-4 |x|: below the bf16 rounding of the
// activation for x > -2 and below 2e-3 absolute everywhere that matters (the reference rounds silu to bf16).
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float silu_f(float x) {
  const float h = 0.5f * x;
  return fmaf(h, tanh_approx(h), h);
}
// d silu / dx = sig * (1 + x * (1 - sig))
__device__ __forceinline__ float dsilu_f(float x) {
  const float sg = fmaf(0.5f, tanh_approx(0.5f * x), 0.5f);
  return sg * fmaf(x, 1.0f - sg, 1.0f);
}

__device__ __forceinline__ void unpack4(uint2 v, float f[4]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
}
__device__ __forceinline__ uint2 pack4(const float f[4]) { return make_uint2(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3])); }
__device__ __forceinline__ void unpack8(const uint4& v, float f[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}

struct GlumbParams {
  const __nv_bfloat16* u;     // [B, H, W, 2*Hd]
  const __nv_bfloat16* w;     // [2*Hd, 9]
  const __nv_bfloat16* bias;  // [2*Hd]
  const __nv_bfloat16* dmid;  // [B, H, W, Hd]   (dc kernel)
  __nv_bfloat16* mid;         // [B, H, W, Hd]   (forward)
  __nv_bfloat16* dc;          // [B, H, W, 2*Hd] (dc kernel out, du kernel in)
  __nv_bfloat16* du;          // [B, H, W, 2*Hd]
  float* partial;             // [B, 2*Hd, 10]   (du kernel)
  int B, H, W, Hd;
};

// One ring row = [kTW + 2 columns][128 channels] fp32; ring column j <-> image column x0 - 1 + j.
// Row loader: 160 of the 256 threads own one 128-bit vector (8 bf16 channels) of every row.  The vector
// travels global -> raw staging ring (cp.async, kDepth rows ahead, no registers held) -> [silu] -> fp32 ring.
// Only the issuing thread reads its staged vector back, so cp.async.wait_group is all the ordering it needs.
constexpr int kDepth = 8;                                     // rows in flight per CTA

struct RowLoader {
  const __nv_bfloat16* src;   // sample base
  int64_t pix_stride;         // elements per pixel
  int H, W, x0;
  int j, v;                   // this thread's ring column / vector
  bool active, valid;         // valid: inside the image and inside the channel range
  int64_t ch;                 // first of its 8 channels in the source tensor
  uint4* stage;               // this thread's slot in stage row 0; rows are kStageRowVecs apart
  static constexpr int kStageRowVecs = (kTW + 2) * 16;
  __device__ __forceinline__ void issue(int r) {               // always commits a group (possibly empty)
    if (active && valid && r >= 0 && r < H) {
      const uint32_t dst = (uint32_t)__cvta_generic_to_shared(stage + (r % kDepth) * kStageRowVecs);
      const void* g = src + ((int64_t)r * W + (x0 - 1 + j)) * pix_stride + ch;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(g) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  // row r was issued kDepth - 1 groups before the newest one
  template <bool SILU>
  __device__ __forceinline__ void commit(float* ring_row, int r) {
    asm volatile("cp.async.wait_group %0;" ::"n"(kDepth - 1) : "memory");
    if (!active) return;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = 0.f;
    if (valid && r >= 0 && r < H) {
      unpack8(stage[(r % kDepth) * kStageRowVecs], f);
      if (SILU) {
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = silu_f(f[e]);
      }
    }
    float4* d = reinterpret_cast<float4*>(ring_row + j * 128 + v * 8);
    d[0] = make_float4(f[0], f[1], f[2], f[3]);
    d[1] = make_float4(f[4], f[5], f[6], f[7]);
  }
};

// ---------------------------------------------------------------------------------------------
// forward (BWD = false) and dc (+ dw, db) kernel (BWD = true).  Thread values [0,1] = the two A
// channels ca, ca+1; [2,3] = their gates Hd+ca, Hd+ca+1.
// ---------------------------------------------------------------------------------------------
template <bool BWD>
__global__ void __launch_bounds__(kThreads, BWD ? 1 : 2) glumb_mid_kernel(const GlumbParams p) {
  __shared__ __align__(16) float ring[kRing][kRowFloats];
  __shared__ __align__(16) uint4 stage[kDepth][RowLoader::kStageRowVecs];
  __shared__ float red[8][4][33];
  const int W = p.W, H = p.H, Hd = p.Hd;
  const int tiles_x = (W + kTW - 1) / kTW;
  const int b = blockIdx.y / tiles_x, tile = blockIdx.y - b * tiles_x;
  const int x0 = tile * kTW, c0 = blockIdx.x * 64;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t C2 = 2 * (int64_t)Hd;

  RowLoader L;
  L.src = p.u + (int64_t)b * H * W * C2;
  L.pix_stride = C2; L.H = H; L.W = W; L.x0 = x0;
  L.active = tid < (kTW + 2) * 16;
  L.j = tid >> 4; L.v = tid & 15;
  {
    const int cl = c0 + 8 * (L.v & 7);                          // A-half channel index
    L.ch = L.v < 8 ? cl : Hd + cl;
    L.valid = (x0 - 1 + L.j) >= 0 && (x0 - 1 + L.j) < W && cl < Hd;
    L.stage = &stage[0][L.active ? tid : 0];
  }

  const int ca = c0 + 2 * lane;
  const int x = x0 + warp;
  const bool ch_ok = ca < Hd;                                   // Hd is even
  const bool ok = ch_ok && x < W;
  // weights / bias of the CTA's 64 + 64 channels: coalesced into shared memory (the stage ring is still idle),
  // then 36 + 4 values per thread into registers
  float w[9][4], bias[4];
  {
    __nv_bfloat16* wsm = reinterpret_cast<__nv_bfloat16*>(&stage[0][0]);     // [2][64][9] then bias [2][64]
    for (int i = tid; i < 2 * 64 * 10; i += kThreads) {
      const int half = i / 640, r = i - half * 640;                           // r < 576: weight, else bias
      const int cl = r < 576 ? r / 9 : r - 576;
      float val = 0.f;
      if (c0 + cl < Hd) {
        const int64_t cg = (half ? Hd : 0) + c0 + cl;
        val = __bfloat162float(r < 576 ? p.w[cg * 9 + (r - cl * 9)] : p.bias[cg]);
      }
      wsm[i] = __float2bfloat16_rn(val);
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int base = (j >> 1) * 640, cl = 2 * lane + (j & 1);
      bias[j] = __bfloat162float(wsm[base + 576 + cl]);
#pragma unroll
      for (int k = 0; k < 9; ++k) w[k][j] = __bfloat162float(wsm[base + cl * 9 + k]);
    }
    __syncthreads();
  }
  float dw[9][4], db[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    db[j] = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) dw[k][j] = 0.f;
  }
  float win0[3][4], win1[3][4], win2[3][4];
#pragma unroll
  for (int kx = 0; kx < 3; ++kx)
#pragma unroll
    for (int j = 0; j < 4; ++j) win0[kx][j] = win1[kx][j] = win2[kx][j] = 0.f;

  auto load_cols = [&](const float* ring_row, float (&dst)[3][4]) {
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const float* px = ring_row + (warp + kx) * 128 + 2 * lane;
      const float2 a = *reinterpret_cast<const float2*>(px);
      const float2 g = *reinterpret_cast<const float2*>(px + 64);
      dst[kx][0] = a.x; dst[kx][1] = a.y; dst[kx][2] = g.x; dst[kx][3] = g.y;
    }
  };

  // prologue: rows 0 .. kDepth-1 in flight; rows 0 and 1 activated into the ring
#pragma unroll
  for (int r = 0; r < kDepth; ++r) L.issue(r);
  L.commit<true>(ring[0], 0);
  L.issue(kDepth);
  L.commit<true>(ring[1], 1);
  __syncthreads();
  load_cols(ring[0], win1);                        // row 0; row -1 (win0) is zero

  // dc kernel: dmid of this thread's pixel, fetched two rows ahead of its use
  const __nv_bfloat16* dm_ptr = BWD && ok ? p.dmid + (((int64_t)b * H) * W + x) * Hd + ca : nullptr;
  const int64_t dm_row = (int64_t)W * Hd;
  uint32_t dm_a = 0, dm_b = 0;
  if (BWD && ok) {
    dm_a = __ldg(reinterpret_cast<const uint32_t*>(dm_ptr));
    if (H > 1) dm_b = __ldg(reinterpret_cast<const uint32_t*>(dm_ptr + dm_row));
  }

  // rows (y-1, y) are in (ra, rb); row y+1 comes out of the ring into rc
  auto step = [&](int y, float (&ra)[3][4], float (&rb)[3][4], float (&rc)[3][4]) {
    L.issue(y + 1 + kDepth);                       // its stage slot held row y+1, activated during step y-1
    uint32_t dmr = dm_a;
    if (BWD) {
      dm_a = dm_b;
      if (ok && y + 2 < H) dm_b = __ldg(reinterpret_cast<const uint32_t*>(dm_ptr + (y + 2) * dm_row));
    }
    load_cols(ring[(y + 1) % kRing], rc);
    float acc[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float a = bias[j];
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        a = fmaf(w[kx][j], ra[kx][j], a);
        a = fmaf(w[3 + kx][j], rb[kx][j], a);
        a = fmaf(w[6 + kx][j], rc[kx][j], a);
      }
      acc[j] = a;
    }
    const int64_t pix = ((int64_t)b * H + y) * W + x;
    if (!BWD) {
      if (ok)
        *reinterpret_cast<uint32_t*>(p.mid + pix * Hd + ca) = pack_bf16(acc[0] * silu_f(acc[2]), acc[1] * silu_f(acc[3]));
    } else if (ok) {
      const float dm0 = bf16_lo(dmr), dm1 = bf16_hi(dmr);
      const uint32_t da = pack_bf16(dm0 * silu_f(acc[2]), dm1 * silu_f(acc[3]));
      const uint32_t dg = pack_bf16(dm0 * acc[0] * dsilu_f(acc[2]), dm1 * acc[1] * dsilu_f(acc[3]));
      *reinterpret_cast<uint32_t*>(p.dc + pix * C2 + ca) = da;
      *reinterpret_cast<uint32_t*>(p.dc + pix * C2 + Hd + ca) = dg;
      const float d[4] = {bf16_lo(da), bf16_hi(da), bf16_lo(dg), bf16_hi(dg)};   // what the du kernel will read
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        db[j] += d[j];
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          dw[kx][j] = fmaf(ra[kx][j], d[j], dw[kx][j]);
          dw[3 + kx][j] = fmaf(rb[kx][j], d[j], dw[3 + kx][j]);
          dw[6 + kx][j] = fmaf(rc[kx][j], d[j], dw[6 + kx][j]);
        }
      }
    }
    L.commit<true>(ring[(y + 2) % kRing], y + 2);
    __syncthreads();
  };
  for (int y0 = 0; y0 < H; y0 += 3) {
    step(y0, win0, win1, win2);
    if (y0 + 1 < H) step(y0 + 1, win1, win2, win0);
    if (y0 + 2 < H) step(y0 + 2, win2, win0, win1);
  }

  if (BWD) {
    // per-CTA sums over the 8 columns, four values at a time (fixed order), -> partial[(b, tile)][channel][10]
    float* part = p.partial + ((int64_t)b * tiles_x + tile) * C2 * 10;
#pragma unroll
    for (int g = 0; g < 10; ++g) {
#pragma unroll
      for (int j = 0; j < 4; ++j) red[warp][j][lane] = g < 9 ? dw[g][j] : db[j];
      __syncthreads();
      if (tid < 128) {
        const int j = tid >> 5, l = tid & 31;
        float s = 0.f;
#pragma unroll
        for (int wv = 0; wv < 8; ++wv) s += red[wv][j][l];
        const int cq = c0 + 2 * l;
        if (cq < Hd) part[((j < 2 ? cq : Hd + cq - 2) + j) * 10 + g] = s;
      }
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------------------
// du kernel: one CTA = (sample b, 8 columns, 128 of the 2*Hd channels); ring of raw dc rows.
//   ds[q] = sum_k w[k] * dc[q - k + 1]        du = ds * silu'(u)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 2) glumb_du_kernel(const GlumbParams p) {
  __shared__ __align__(16) float ring[kRing][kRowFloats];
  __shared__ __align__(16) uint4 stage[kDepth][RowLoader::kStageRowVecs];
  const int W = p.W, H = p.H;
  const int64_t C2 = 2 * (int64_t)p.Hd;
  const int tiles_x = (W + kTW - 1) / kTW;
  const int b = blockIdx.y / tiles_x, tile = blockIdx.y - b * tiles_x;
  const int x0 = tile * kTW, c0 = blockIdx.x * 128;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  RowLoader L;
  L.src = p.dc + (int64_t)b * H * W * C2;
  L.pix_stride = C2; L.H = H; L.W = W; L.x0 = x0;
  L.active = tid < (kTW + 2) * 16;
  L.j = tid >> 4; L.v = tid & 15;
  L.ch = c0 + 8 * L.v;
  L.valid = (x0 - 1 + L.j) >= 0 && (x0 - 1 + L.j) < W && L.ch < C2;
  L.stage = &stage[0][L.active ? tid : 0];

  const int64_t ch = c0 + 4 * lane;
  const int x = x0 + warp;
  const bool ok = ch < C2 && x < W;
  const __nv_bfloat16* ub = p.u + (int64_t)b * H * W * C2;
  __nv_bfloat16* dub = p.du + (int64_t)b * H * W * C2;
  float w[9][4];
  {
    __nv_bfloat16* wsm = reinterpret_cast<__nv_bfloat16*>(&stage[0][0]);     // [128][9], coalesced
    for (int i = tid; i < 128 * 9; i += kThreads)
      wsm[i] = (c0 + i / 9) < C2 ? p.w[(int64_t)c0 * 9 + i] : __float2bfloat16_rn(0.f);
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int k = 0; k < 9; ++k) w[k][j] = __bfloat162float(wsm[(4 * lane + j) * 9 + k]);
    __syncthreads();
  }

  float win0[3][4], win1[3][4], win2[3][4];
#pragma unroll
  for (int kx = 0; kx < 3; ++kx)
#pragma unroll
    for (int j = 0; j < 4; ++j) win0[kx][j] = win1[kx][j] = win2[kx][j] = 0.f;

  auto load_cols = [&](const float* ring_row, float (&dst)[3][4]) {
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const float4 d = *reinterpret_cast<const float4*>(ring_row + (warp + kx) * 128 + 4 * lane);
      dst[kx][0] = d.x; dst[kx][1] = d.y; dst[kx][2] = d.z; dst[kx][3] = d.w;
    }
  };

#pragma unroll
  for (int r = 0; r < kDepth; ++r) L.issue(r);
  L.commit<false>(ring[0], 0);
  L.issue(kDepth);
  L.commit<false>(ring[1], 1);
  __syncthreads();
  load_cols(ring[0], win1);

  // u of this thread's pixel, fetched two rows ahead of its use
  const __nv_bfloat16* u_ptr = ub + (int64_t)x * C2 + ch;
  const int64_t u_row = (int64_t)W * C2;
  uint2 u_a = make_uint2(0, 0), u_b = make_uint2(0, 0);
  if (ok) {
    u_a = __ldg(reinterpret_cast<const uint2*>(u_ptr));
    if (H > 1) u_b = __ldg(reinterpret_cast<const uint2*>(u_ptr + u_row));
  }

  // dc rows (y-1, y) are in (ra, rb); row y+1 comes out of the ring into rc.
  // tap (ky, kx) pairs with dc at (y - ky + 1, x - kx + 1): window row 2-ky, window column 2-kx.
  auto step = [&](int y, float (&ra)[3][4], float (&rb)[3][4], float (&rc)[3][4]) {
    L.issue(y + 1 + kDepth);
    const int64_t pix = (int64_t)y * W + x;
    const uint2 uraw = u_a;
    u_a = u_b;
    if (ok && y + 2 < H) u_b = __ldg(reinterpret_cast<const uint2*>(u_ptr + (y + 2) * u_row));
    load_cols(ring[(y + 1) % kRing], rc);
    if (ok) {
      float uu[4], o[4];
      unpack4(uraw, uu);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float a = 0.f;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          a = fmaf(w[kx][j], rc[2 - kx][j], a);
          a = fmaf(w[3 + kx][j], rb[2 - kx][j], a);
          a = fmaf(w[6 + kx][j], ra[2 - kx][j], a);
        }
        o[j] = a * dsilu_f(uu[j]);
      }
      *reinterpret_cast<uint2*>(dub + pix * C2 + ch) = pack4(o);
    }
    L.commit<false>(ring[(y + 2) % kRing], y + 2);
    __syncthreads();
  };
  for (int y0 = 0; y0 < H; y0 += 3) {
    step(y0, win0, win1, win2);
    if (y0 + 1 < H) step(y0 + 1, win1, win2, win0);
    if (y0 + 2 < H) step(y0 + 2, win2, win0, win1);
  }
}

__global__ void __launch_bounds__(256) glumb_finish_kernel(const float* __restrict__ partial, int parts, int64_t n10,
                                                           __nv_bfloat16* __restrict__ dw, __nv_bfloat16* __restrict__ db) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;   // over [2*Hd][10]
  if (i >= n10) return;
  float s = 0.f;
  for (int b = 0; b < parts; ++b) s += partial[(int64_t)b * n10 + i];
  const int64_t ch = i / 10;
  const int k = (int)(i - ch * 10);
  if (k < 9) dw[ch * 9 + k] = __float2bfloat16_rn(s);
  else       db[ch] = __float2bfloat16_rn(s);
}

inline bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; }
inline int tiles_of(int W) { return (W + kTW - 1) / kTW; }

int check_shape(const char* what, int B, int H, int W, int Hd) {
  YAT_REQUIRE(B > 0 && H > 0 && W > 0 && Hd > 0, YAT_ERR_BAD_ARG, "%s: B=%d H=%d W=%d Hd=%d must be positive", what, B, H, W, Hd);
  YAT_REQUIRE(Hd % 8 == 0, YAT_ERR_UNSUPPORTED, "%s: hidden width %d must be a multiple of 8", what, Hd);
  YAT_REQUIRE((int64_t)B * tiles_of(W) < 65536, YAT_ERR_UNSUPPORTED, "%s: B * x-tiles = %lld exceeds the grid limit", what,
              (long long)B * tiles_of(W));
  return YAT_OK;
}
}  // namespace
}  // namespace yat

extern "C" int64_t yat_glumb_workspace_bytes(int B, int W, int Hd) {
  return (int64_t)B * yat::tiles_of(W) * 2 * Hd * 10 * (int64_t)sizeof(float);
}

extern "C" int yat_glumb_mid_fwd(const void* u, const void* w_dw, const void* b_dw, int B, int H, int W, int Hd,
                                 void* mid, void* stream) {
  using namespace yat;
  YAT_REQUIRE(u && w_dw && b_dw && mid, YAT_ERR_BAD_ARG, "yat_glumb_mid_fwd: null pointer argument");
  int rc = check_shape("yat_glumb_mid_fwd", B, H, W, Hd);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(u) && aligned16(mid), YAT_ERR_BAD_ARG, "yat_glumb_mid_fwd: tensors must be 16-byte aligned");
  GlumbParams p{};
  p.u = (const __nv_bfloat16*)u; p.w = (const __nv_bfloat16*)w_dw; p.bias = (const __nv_bfloat16*)b_dw;
  p.mid = (__nv_bfloat16*)mid; p.B = B; p.H = H; p.W = W; p.Hd = Hd;
  glumb_mid_kernel<false><<<dim3((Hd + 63) / 64, B * tiles_of(W)), kThreads, 0, (cudaStream_t)stream>>>(p);
  return check_launch("glumb_mid_fwd");
}

extern "C" int yat_glumb_mid_bwd(const void* u, const void* w_dw, const void* b_dw, const void* dmid, int B, int H,
                                 int W, int Hd, void* dc, void* du, void* dw, void* db, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(u && w_dw && b_dw && dmid && dc && du && dw && db && workspace, YAT_ERR_BAD_ARG,
              "yat_glumb_mid_bwd: null pointer argument");
  int rc = check_shape("yat_glumb_mid_bwd", B, H, W, Hd);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(u) && aligned16(dmid) && aligned16(dc) && aligned16(du), YAT_ERR_BAD_ARG,
              "yat_glumb_mid_bwd: tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  GlumbParams p{};
  p.u = (const __nv_bfloat16*)u; p.w = (const __nv_bfloat16*)w_dw; p.bias = (const __nv_bfloat16*)b_dw;
  p.dmid = (const __nv_bfloat16*)dmid; p.dc = (__nv_bfloat16*)dc; p.du = (__nv_bfloat16*)du;
  p.partial = (float*)workspace; p.B = B; p.H = H; p.W = W; p.Hd = Hd;
  const int parts = B * tiles_of(W);
  glumb_mid_kernel<true><<<dim3((Hd + 63) / 64, parts), kThreads, 0, st>>>(p);
  if ((rc = check_launch("glumb_dc")) != YAT_OK) return rc;
  glumb_du_kernel<<<dim3((2 * Hd + 127) / 128, parts), kThreads, 0, st>>>(p);
  if ((rc = check_launch("glumb_du")) != YAT_OK) return rc;
  const int64_t n10 = (int64_t)2 * Hd * 10;
  glumb_finish_kernel<<<(unsigned)((n10 + 255) / 256), 256, 0, st>>>(p.partial, parts, n10, (__nv_bfloat16*)dw,
                                                                   (__nv_bfloat16*)db);
  return check_launch("glumb_finish");
}
