// Fused middle of SANA's GLUMBConv feed-forward (the reference runs it as
// nn.SiLU + cuDNN depthwise conv + bias + chunk + SiLU + mul, utils/patched_sana_transformer.py GLUMBConv
// via utils/patch_sana_attention_layers.py:106-111):
//
//   u   [B, H, W, 2*Hd] bf16   output of conv_inverted (1x1, run as a GEMM on token-major data)
//   s   = silu(u)                                   (kept in fp32; the reference rounds it to bf16)
//   c   = depthwise3x3(s; w[2*Hd][3][3]) + b        (zero padding 1, fp32 accumulation)
//   mid = c[:Hd] * silu(c[Hd:])        [B, H, W, Hd] bf16   -> input of conv_point (1x1 GEMM)
//
// Token-major (= NHWC) throughout: channels are the contiguous axis, so every warp access is one full
// 128-byte line (32 lanes x one bf16 channel pair) and there are no NCHW<->NHWC permutes.
//
// Register-strip design, no CTA barriers in the sweep: lane = channel pair, warp = a strip of kR = 4 image
// rows, CTA = 8 strips (32 rows) x 64 channels, sweeping x = 0 .. W-1.  Each step consumes ONE image column of
// the CTA (32 + 2 rows with the halo), delivered by TMA: a ninth (producer) warp issues one
// cp.async.bulk.tensor box {64 channels, 1 column, 34 rows} per tensor into a CTA-wide 4-slot shared-memory
// ring (full / empty mbarriers per slot, out-of-image rows / columns zero-filled by the TMA unit), so the
// compute warps spend no instructions on addresses, bounds, zero padding or tensor-map issue and hold no
// registers for loads in flight.  SiLU is applied once per loaded element and the column is scattered into
// three sets of fp32 accumulators - the outputs at x-1, x and x+1 - so a conv output costs its 9 FMAs and
// nothing else is kept.  Each warp reads its 4 + 2 rows from the shared column (the halo rows between strips
// are fetched once).  History (ncu, SANA-1.6B layer): shared-memory ring with a CTA barrier per row 222 M warp
// instructions / 345 us; register strip with per-thread global loads 186 M / 240 us; per-warp TMA rings
// (every warp issuing its own boxes) 140 M / 217 us.
//
// Backward (same sweep; CTA = one half, A or gate; warp = row strip): the forward also stores c = conv output
// (bf16, what the reference keeps for autograd anyway); every step forms one column of
// dc = [dmid * silu(g) | dmid * xa * silu'(g)] in registers (dc never touches HBM), keeps a 3-column window of
// it, and computes du = (depthwise3x3^T dc) * silu'(u) together with dw[k] += silu(u) * dc[tap k], db += dc;
// per-CTA partial sums -> fixed-order finish (deterministic).
#include "common.cuh"
#include "sm100.cuh"

namespace yat {
namespace {

using namespace sm100;

constexpr int kCW = 8;                             // consumer warps = row strips per CTA (8 -> 32 rows)
constexpr int kThreads = kCW * 32;
constexpr int kR = 4;                              // output rows per warp
constexpr int kIn = kR + 2;                        // input rows per warp and column (halo above and below)
constexpr int kRowsCta = kCW * kR;                 // 32 output rows per CTA
constexpr int kInCta = kRowsCta + 2;               // 34 input rows per CTA and column: the halo rows between strips are fetched ONCE
constexpr int kPart = 11;                          // backward partials per channel: dw[9], db, sum(du)
constexpr int kStages = 6;                         // ring slots per CTA (two groups of three: the sweep is unrolled by 3)
constexpr int kAhead = 3;                          // column t+3 is issued in step t into the slot read in step t-3
constexpr int kRowB = 128;                         // bytes of one box row: 64 bf16 channels
constexpr int kFwdSlotB = 2 * kInCta * kRowB;                      // u: A rows | gate rows
constexpr int kBwdSlotB0 = (2 * kInCta + kRowsCta) * kRowB;        // half A:  g | dmid | u
constexpr int kBwdSlotB1 = (3 * kInCta + kRowsCta) * kRowB;        // half G:  g | dmid | u | xa
constexpr int kBarBytes = 2 * kStages * 8;                         // full[kStages], empty[kStages]
constexpr int kFwdSmem = kStages * kFwdSlotB + kBarBytes;
constexpr int kBwdSmem = kStages * kBwdSlotB1 + kBarBytes;

// silu through ONE special-function op:  sigmoid(x) = 0.5 + 0.5 tanh(x / 2)  =>  silu(x) = h + h tanh(h), h = x / 2.
// tanh.approx.f32 has 2^-11 relative error, i.e. |error(silu)| <= 2.5e-4 |x|: below the bf16 rounding of the
// activation for x > -2 and below 2e-3 absolute everywhere that matters (the reference rounds silu to bf16).
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ f2 tanh2(f2 h) {
  float a, b;
  un2(h, a, b);
  return mk2(tanh_approx(a), tanh_approx(b));
}
__device__ __forceinline__ f2 silu2(f2 x) {
  const f2 h = mul2(x, bc2(0.5f));
  return fma2(h, tanh2(h), h);
}
// s = silu(x), d = silu'(x) = sg (1 + x (1 - sg))
__device__ __forceinline__ f2 sigmoid2(f2 x) { return fma2(bc2(0.5f), tanh2(mul2(x, bc2(0.5f))), bc2(0.5f)); }
__device__ __forceinline__ f2 silu_grad_from_sigmoid2(f2 x, f2 sg) {
  const f2 one = bc2(1.0f);
  return mul2(sg, fma2(x, fma2(sg, bc2(-1.0f), one), one));
}
// With h = x / 2, th = tanh(h):  silu(x) = h + h th  and  2 silu'(x) = 1 + r,  r = th + h (1 - th^2).  The backward
// works from nh = -h (one multiply by -0.5; the special-function op takes the negated source for free), so that no
// operand ever needs a separate negation:  -silu(x) = nh + nh th,  r = th + nh (th^2 - 1).  Signs and the factors
// 1/2 are carried in the conv weights and in the final scale of dw / db (powers of two: exact).
__device__ __forceinline__ f2 tanh2_neg(f2 nh) {                 // tanh(-nh)
  float a, b;
  un2(nh, a, b);
  return mk2(tanh_approx(-a), tanh_approx(-b));
}
__device__ __forceinline__ f2 silu_grad_r(f2 nh, f2 th) { return fma2(nh, fma2(th, th, bc2(-1.0f)), th); }
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t r;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(r) : "r"(addr));
  return r;
}

struct GlumbParams {
  const __nv_bfloat16* w;     // [2*Hd, 9]
  const __nv_bfloat16* bias;  // [2*Hd]
  __nv_bfloat16* mid;         // [B, H, W, Hd]   (forward out)
  __nv_bfloat16* c;           // [B, H, W, 2*Hd] conv output (forward out, optional)
  __nv_bfloat16* du;          // [B, H, W, 2*Hd] (backward out)
  float* partial;             // [B * y-tiles, 2*Hd, kPart] (backward)
  int B, H, W, Hd;
};
struct alignas(64) GlumbArgs {
  CUtensorMap tm_u;           // u    {2*Hd, W, H, B}, box {64, 1, kInCta, 1} (forward) / {64, 1, kRowsCta, 1} (backward)
  CUtensorMap tm_c;           // c    {2*Hd, W, H, B}, box {64, 1, kInCta, 1}                          (backward)
  CUtensorMap tm_d;           // dmid {Hd,   W, H, B}, box {64, 1, kInCta, 1}                          (backward)
  GlumbParams p;
};

inline int ytiles_of(int H) { return (H + kRowsCta - 1) / kRowsCta; }

// Ring protocol (both kernels).  Column t lives in slot t % 6 = 3 (m & 1) + j for t = 3 m + j (the sweep is unrolled
// by three).  In step t ONE warp (the live warps take turns) issues the CTA-wide boxes of column t + 3: its lane 0
// waits for empty[slot] (every live warp has read column t - 3: two steps of slack, so the wait rarely spins),
// posts the byte count on full[slot] and issues the loads; all coordinates are CTA-uniform, so the issue path is
// a handful of uniform-datapath instructions executed once per CTA and column.  (The previous version issued
// per-warp boxes from every warp in every step: ~75 of its ~340 instructions per step were the tensor-map issue
// sequence, and the halo rows between strips were fetched twice.)  Consumers: wait full[slot], read their kIn
// rows, __syncwarp, lane 0 arrives on empty[slot].  Every waiter observes every phase of its barrier in order.
struct Ring {
  uint32_t base, full, empty;
};
__device__ __forceinline__ Ring ring_setup(uint8_t* smem, int slot_bytes, int live_warps) {
  Ring r;
  r.base = smem_u32(smem);
  r.full = r.base + kStages * slot_bytes;
  r.empty = r.full + kStages * 8;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) {
      mbar_init(r.full + 8 * s, 1);
      mbar_init(r.empty + 8 * s, live_warps);
    }
    fence_barrier_init();
  }
  __syncthreads();
  return r;
}

// ---------------------------------------------------------------------------------------------
// forward.  Thread values [0] = the two A channels ca, ca+1; [1] = their gates Hd+ca, Hd+ca+1.
// Optionally writes c (the depthwise-conv output, bf16) for the backward.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 2) glumb_fwd_kernel(const __grid_constant__ GlumbArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const GlumbParams& p = a.p;
  const int W = p.W, H = p.H, Hd = p.Hd;
  const int ytiles = (H + kRowsCta - 1) / kRowsCta;
  const int b = blockIdx.y / ytiles, yt = blockIdx.y - b * ytiles;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Y0 = yt * kRowsCta;
  const int c0 = blockIdx.x * 64;
  const int live_warps = min(kCW, (H - Y0 + kR - 1) / kR);
  const Ring ring = ring_setup(smem, kFwdSlotB, live_warps);

  if (warp >= live_warps) return;                                  // strip below the image (not counted in empty[])
  // column tt (rows Y0-1 .. Y0+32, A half then gate half) -> slot; wait_phase < 0: the slot has not been used yet
  auto issue = [&](int tt, int slot, int wait_phase) {
    if (lane == 0) {
      if (wait_phase >= 0) mbar_wait(ring.empty + 8 * slot, (uint32_t)wait_phase & 1u);
      const uint32_t bar = ring.full + 8 * slot, dst = ring.base + slot * kFwdSlotB;
      mbar_arrive_expect_tx(bar, kFwdSlotB);
      tma_load_4d(dst, &a.tm_u, bar, c0, tt, Y0 - 1, b);
      tma_load_4d(dst + kInCta * kRowB, &a.tm_u, bar, Hd + c0, tt, Y0 - 1, b);
    }
  };
  if (warp == 0) {
#pragma unroll
    for (int t = 0; t < kAhead; ++t)
      if (t <= W) issue(t, t, -1);
  }
  int issuer = 0;                                                  // the live warp that issues in the current step
  const int y0 = Y0 + warp * kR;
  const int ca = c0 + 2 * lane;
  const bool live = ca < Hd;                                      // Hd is even
  const int64_t C2 = 2 * (int64_t)Hd;

  // [0] = the A channel pair (ca, ca+1), [1] = their gates (Hd+ca, Hd+ca+1)
  f2 w[9][2], bias[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int64_t ch = j * (int64_t)Hd + ca;
    bias[j] = live ? mk2(__bfloat162float(p.bias[ch]), __bfloat162float(p.bias[ch + 1])) : bc2(0.f);
#pragma unroll
    for (int k = 0; k < 9; ++k)
      w[k][j] = live ? mk2(__bfloat162float(p.w[ch * 9 + k]), __bfloat162float(p.w[(ch + 1) * 9 + k])) : bc2(0.f);
  }

  // outputs: running byte pointers advanced once per step; a row is one 32x32->64 multiply-add away
  const uint32_t col_b = (uint32_t)C2 * 2u, row_b = (uint32_t)W * col_b;          // bytes (checked < 2^32 by the launcher)
  const uint32_t mcol_b = (uint32_t)Hd * 2u, mid_row_b = (uint32_t)W * mcol_b;
  char* out_m = reinterpret_cast<char*>(p.mid + (((int64_t)b * H + y0) * W) * Hd + ca);
  char* out_a = p.c ? reinterpret_cast<char*>(p.c + (((int64_t)b * H + y0) * W) * C2 + ca) : nullptr;
  char* out_g = out_a + mcol_b;
  uint32_t omask = 0;
#pragma unroll
  for (int r = 0; r < kR; ++r) omask |= (live && y0 + r < H) ? (1u << r) : 0u;

  f2 acc0[kR][2], acc1[kR][2], acc2[kR][2];
#pragma unroll
  for (int r = 0; r < kR; ++r)
#pragma unroll
    for (int j = 0; j < 2; ++j) acc0[r][j] = acc1[r][j] = acc2[r][j] = bias[j];

  const uint32_t my_rows = ring.base + (warp * kR) * kRowB + lane * 4;   // this strip's first input row (y0 - 1) in a slot

  // step x: input column x feeds the outputs at x-1 (tap kx = 2: aP, finished and written here), x (kx = 1: aC)
  // and x+1 (kx = 0: aN, started here from the bias).  x = W is the zero column right of the image.
  auto step = [&](int m, int j, int x, f2 (&aP)[kR][2], f2 (&aC)[kR][2], f2 (&aN)[kR][2]) {   // x = 3 m + j
    if (x + kAhead <= W && warp == issuer) issue(x + kAhead, 3 * ((m + 1) & 1) + j, ((m + 1) >> 1) - 1);
    issuer = (issuer + 1 == live_warps) ? 0 : issuer + 1;
    const int s_idx = 3 * (m & 1) + j;
    mbar_wait(ring.full + 8 * s_idx, (uint32_t)(m >> 1) & 1u);
    const uint32_t col = my_rows + s_idx * kFwdSlotB;
    uint32_t raw[2][kIn];
#pragma unroll
    for (int i = 0; i < kIn; ++i) {
      raw[0][i] = lds32(col + i * kRowB);
      raw[1][i] = lds32(col + (kInCta + i) * kRowB);
    }
    __syncwarp();                                                 // every lane has read the slot
    if (lane == 0) mbar_arrive(ring.empty + 8 * s_idx);
#pragma unroll
    for (int j = 0; j < 2; ++j) {                                 // A channels, then their gates: 6 activations live at a time
      f2 s[kIn];
#pragma unroll
      for (int i = 0; i < kIn; ++i) s[i] = silu2(unpack2(raw[j][i]));
#pragma unroll
      for (int r = 0; r < kR; ++r) {
        f2 pa = aP[r][j], cc = aC[r][j], n = bias[j];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          pa = fma2(w[3 * ky + 2][j], s[r + ky], pa);
          cc = fma2(w[3 * ky + 1][j], s[r + ky], cc);
          n = fma2(w[3 * ky + 0][j], s[r + ky], n);
        }
        aP[r][j] = pa; aC[r][j] = cc; aN[r][j] = n;
      }
    }
    if (x > 0) {
#pragma unroll
      for (int r = 0; r < kR; ++r) {
        // c is what the reference keeps as the bf16 conv output; the GLU uses the same rounded values
        const uint32_t cxa = pack2(aP[r][0]), cg = pack2(aP[r][1]);
        const uint32_t m = pack2(mul2(unpack2(cxa), silu2(unpack2(cg))));
        if ((omask >> r) & 1u) {
          if (out_a) {
            *reinterpret_cast<uint32_t*>(out_a + (uint64_t)row_b * (uint32_t)r) = cxa;
            *reinterpret_cast<uint32_t*>(out_g + (uint64_t)row_b * (uint32_t)r) = cg;
          }
          *reinterpret_cast<uint32_t*>(out_m + (uint64_t)mid_row_b * (uint32_t)r) = m;
        }
      }
      out_m += mcol_b;
      if (out_a) { out_a += col_b; out_g += col_b; }
    }
  };
  for (int m = 0, x0 = 0; x0 <= W; ++m, x0 += 3) {
    step(m, 0, x0, acc0, acc1, acc2);
    if (x0 + 1 <= W) step(m, 1, x0 + 1, acc1, acc2, acc0);
    if (x0 + 2 <= W) step(m, 2, x0 + 2, acc2, acc0, acc1);
  }
}

// ---------------------------------------------------------------------------------------------
// backward, ONE kernel.  CTA = (64 channels of one half, 8 row strips): half 0 owns the A channels (xa), half 1
// the gates (lowest bit of blockIdx.x, so the two CTAs that read the same g / dmid columns are launched next to
// each other and the second read hits L2); warp = row strip, lane l owns channels 2l, 2l+1 of its half.  With
// c = [xa | g] saved by the forward,
//   dc_A = dmid * silu(g)        dc_G = dmid * xa * silu'(g)
//   ds[q]  = sum_k w[k] * dc[q - k + 1]        du = ds * silu'(u)
//   dw[k] += silu(u)[q] * dc[q - k + 1]        db += dc[q]
// Step t forms dc column t (kIn rows) into the newest slot of a 3-column register window and finishes the
// outputs of column t-1; ring slot t holds the dc inputs of column t and u of column t-1.  All arithmetic is
// unconditional (rows below the image and dead channels see zero-filled inputs / zero weights); only the du
// store is predicated, so the four rows' FMA chains sit in one basic block and interleave.
// dw / db: per-CTA sums over its strips in fixed order -> partial[(b, y-tile)][ch][11].
// ---------------------------------------------------------------------------------------------
template <int KIND>
__device__ __forceinline__ void glumb_bwd_body(const GlumbArgs& a, uint8_t* smem) {
  constexpr int kSlotB = KIND ? kBwdSlotB1 : kBwdSlotB0;
  const GlumbParams& p = a.p;
  const int W = p.W, H = p.H, Hd = p.Hd;
  const int ytiles = (H + kRowsCta - 1) / kRowsCta;
  const int b = blockIdx.y / ytiles, yt = blockIdx.y - b * ytiles;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int Y0 = yt * kRowsCta;
  const int c0 = (blockIdx.x >> 1) * 64, cl = c0 + 2 * lane;
  const bool live = cl < Hd;
  const int64_t C2 = 2 * (int64_t)Hd;
  const int64_t ch = (KIND ? Hd : 0) + cl;
  const int live_warps = min(kCW, (H - Y0 + kR - 1) / kR);
  const Ring ring = ring_setup(smem, kSlotB, live_warps);

  f2 dw[9], db = bc2(0.f), dbu = bc2(0.f);
#pragma unroll
  for (int k = 0; k < 9; ++k) dw[k] = bc2(0.f);

  if (warp < live_warps) {
    const int y0 = Y0 + warp * kR;
    f2 w[9];
#pragma unroll
    for (int k = 0; k < 9; ++k)
      w[k] = live ? mul2(bc2(KIND ? 0.25f : -0.5f), mk2(__bfloat162float(p.w[ch * 9 + k]), __bfloat162float(p.w[(ch + 1) * 9 + k])))
                  : bc2(0.f);         // x 1/2: silu'(u) = (1 + r) / 2;  A half carries -dc (x -1), gate half 2 dc (x 1/2)

    const uint32_t col_b = (uint32_t)C2 * 2u, row_b = (uint32_t)W * col_b;
    char* out_u = reinterpret_cast<char*>(p.du + (((int64_t)b * H + y0) * W) * C2 + ch);
    uint32_t omask = 0;
#pragma unroll
    for (int r = 0; r < kR; ++r) omask |= (live && y0 + r < H) ? (1u << r) : 0u;

    f2 win0[kIn], win1[kIn], win2[kIn];
#pragma unroll
    for (int i = 0; i < kIn; ++i) win0[i] = win1[i] = win2[i] = bc2(0.f);

    const uint32_t my_rows = ring.base + (warp * kR) * kRowB + lane * 4;   // this strip's halo row y0 - 1 of the g block

    // slot <- g, dmid (and xa) of column tt, rows Y0-1 .. Y0+32; u of column tt-1, rows Y0 .. Y0+31
    auto issue = [&](int tt, int slot, int wait_phase) {
      if (lane == 0) {
        if (wait_phase >= 0) mbar_wait(ring.empty + 8 * slot, (uint32_t)wait_phase & 1u);
        const uint32_t bar = ring.full + 8 * slot, dst = ring.base + slot * kSlotB;
        mbar_arrive_expect_tx(bar, kSlotB);
        tma_load_4d(dst, &a.tm_c, bar, Hd + c0, tt, Y0 - 1, b);
        tma_load_4d(dst + kInCta * kRowB, &a.tm_d, bar, c0, tt, Y0 - 1, b);
        tma_load_4d(dst + 2 * kInCta * kRowB, &a.tm_u, bar, (KIND ? Hd : 0) + c0, tt - 1, Y0, b);
        if (KIND) tma_load_4d(dst + (2 * kInCta + kRowsCta) * kRowB, &a.tm_c, bar, c0, tt, Y0 - 1, b);
      }
    };
    if (warp == 0) {
#pragma unroll
      for (int t = 0; t < kAhead; ++t)
        if (t <= W) issue(t, t, -1);
    }
    int issuer = 0;

    // step t: dc column t -> cP; outputs of column t-1 from (cM, cC, cP) = dc columns (t-2, t-1, t).
    // tap (ky, kx) pairs with dc at (y - ky + 1, x - kx + 1): window row r - ky + 2, column kx = 0 -> cP, 1 -> cC, 2 -> cM.
    auto step = [&](int m, int j, int t, f2 (&cM)[kIn], f2 (&cC)[kIn], f2 (&cP)[kIn]) {        // t = 3 m + j
      if (t + kAhead <= W && warp == issuer) issue(t + kAhead, 3 * ((m + 1) & 1) + j, ((m + 1) >> 1) - 1);
      issuer = (issuer + 1 == live_warps) ? 0 : issuer + 1;
      const int s_idx = 3 * (m & 1) + j;
      mbar_wait(ring.full + 8 * s_idx, (uint32_t)(m >> 1) & 1u);
      const uint32_t col = my_rows + s_idx * kSlotB;
#pragma unroll
      for (int i = 0; i < kIn; ++i) {
        const f2 g = unpack2(lds32(col + i * kRowB)), dm = unpack2(lds32(col + (kInCta + i) * kRowB));
        const f2 nh = mul2(g, bc2(-0.5f)), th = tanh2_neg(nh);
        if (KIND) {
          const f2 dmxa = mul2(dm, unpack2(lds32(col + (2 * kInCta + kRowsCta + i) * kRowB)));
          cP[i] = fma2(dmxa, silu_grad_r(nh, th), dmxa);                                 // 2 dmid * xa * silu'(g)
        } else {
          cP[i] = mul2(dm, fma2(nh, th, nh));                                            // -dmid * silu(g)
        }
      }
      uint32_t ru[kR];
#pragma unroll
      for (int r = 0; r < kR; ++r) ru[r] = lds32(col + (2 * kInCta + r) * kRowB);        // u block starts at row Y0
      __syncwarp();                                               // every lane has read the slot
      if (lane == 0) mbar_arrive(ring.empty + 8 * s_idx);
      if (t > 0) {
        uint32_t o[kR];
#pragma unroll
        for (int r = 0; r < kR; ++r) {
          const f2 nh = mul2(unpack2(ru[r]), bc2(-0.5f)), th = tanh2_neg(nh);
          const f2 sv = fma2(nh, th, nh), rr = silu_grad_r(nh, th);   // -silu(u); 2 silu'(u) - 1
          f2 acc;                                                 // the four rows' chains interleave (one basic block)
#pragma unroll
          for (int ky = 0; ky < 3; ++ky) {
            const f2 d0 = cP[r - ky + 2], d1 = cC[r - ky + 2], d2 = cM[r - ky + 2];
            acc = ky ? fma2(w[3 * ky + 0], d0, acc) : mul2(w[0], d0);
            acc = fma2(w[3 * ky + 1], d1, acc);
            acc = fma2(w[3 * ky + 2], d2, acc);
            dw[3 * ky + 0] = fma2(sv, d0, dw[3 * ky + 0]);
            dw[3 * ky + 1] = fma2(sv, d1, dw[3 * ky + 1]);
            dw[3 * ky + 2] = fma2(sv, d2, dw[3 * ky + 2]);
          }
          db = add2(db, cC[r + 1]);
          const f2 ov = fma2(acc, rr, acc);                       // (w/2 (*) dc) (1 + r) = ds * silu'(u)
          // dw accumulates (-silu(u)) * (-dc) (A half) or (-silu(u)) * (2 dc) (gate half, final scale -1/2)
          if ((omask >> r) & 1u) dbu = add2(dbu, ov);             // column sum of du = bias gradient of conv_inverted
                                                                  // (a row below the image still sees dc of the last row)
          o[r] = pack2(ov);
        }
#pragma unroll
        for (int r = 0; r < kR; ++r)
          if ((omask >> r) & 1u) *reinterpret_cast<uint32_t*>(out_u + (uint64_t)row_b * (uint32_t)r) = o[r];
        out_u += col_b;
      }
    };
    for (int m = 0, t0 = 0; t0 <= W; ++m, t0 += 3) {
      step(m, 0, t0, win0, win1, win2);
      if (t0 + 1 <= W) step(m, 1, t0 + 1, win1, win2, win0);
      if (t0 + 2 <= W) step(m, 2, t0 + 2, win2, win0, win1);
    }
  }

  // per-CTA sums over the strips, fixed order.  The scratch aliases the ring: every TMA write has been waited
  // for and every slot read is done once all warps have passed the barrier.
  __syncthreads();
  float (*red)[2 * kPart][32] = reinterpret_cast<float (*)[2 * kPart][32]>(smem);
  {
#pragma unroll
    for (int g = 0; g < kPart; ++g) {
      float lo, hi;
      un2(g < 9 ? dw[g] : (g == 9 ? db : dbu), lo, hi);
      const float fs = KIND ? (g < 9 ? -0.5f : (g == 9 ? 0.5f : 1.f)) : (g == 9 ? -1.f : 1.f);   // see the sign / scale notes above
      lo *= fs; hi *= fs;
      red[warp][2 * g + 0][lane] = live ? lo : 0.f;
      red[warp][2 * g + 1][lane] = live ? hi : 0.f;
    }
  }
  __syncthreads();
  float* part = p.partial + ((int64_t)b * ytiles + yt) * C2 * kPart;
  for (int i = tid; i < 2 * kPart * 32; i += kThreads) {           // (j, g, lane)
    const int l = i & 31, g = (i >> 5) % kPart, j = i / (32 * kPart);
    float s = 0.f;
#pragma unroll
    for (int st = 0; st < kCW; ++st) s += red[st][2 * g + j][l];
    const int cq = c0 + 2 * l;
    if (cq < Hd) part[((KIND ? Hd : 0) + cq + j) * kPart + g] = s;
  }
}

// blockIdx.x & 1 = half: 0 -> A channels, 1 -> gate channels (separate instantiations: the A half never loads xa)
__global__ void __launch_bounds__(kThreads, 2) glumb_bwd_kernel(const __grid_constant__ GlumbArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  if (blockIdx.x & 1) glumb_bwd_body<1>(a, smem);
  else                glumb_bwd_body<0>(a, smem);
}

__global__ void __launch_bounds__(256) glumb_finish_kernel(const float* __restrict__ partial, int parts, int64_t n,
                                                           __nv_bfloat16* __restrict__ dw, __nv_bfloat16* __restrict__ db,
                                                           __nv_bfloat16* __restrict__ dbu) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;   // over [2*Hd][kPart]
  if (i >= n) return;
  float s = 0.f;
  for (int b = 0; b < parts; ++b) s += partial[(int64_t)b * n + i];
  const int64_t ch = i / kPart;
  const int k = (int)(i - ch * kPart);
  if (k < 9)        dw[ch * 9 + k] = __float2bfloat16_rn(s);
  else if (k == 9)  db[ch] = __float2bfloat16_rn(s);
  else if (dbu)     dbu[ch] = __float2bfloat16_rn(s);
}

inline bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; }

int check_shape(const char* what, int B, int H, int W, int Hd) {
  YAT_REQUIRE(B > 0 && H > 0 && W > 0 && Hd > 0, YAT_ERR_BAD_ARG, "%s: B=%d H=%d W=%d Hd=%d must be positive", what, B, H, W, Hd);
  YAT_REQUIRE(Hd % 8 == 0, YAT_ERR_UNSUPPORTED, "%s: hidden width %d must be a multiple of 8", what, Hd);
  YAT_REQUIRE((int64_t)W * Hd * 4 < (1ll << 32), YAT_ERR_UNSUPPORTED, "%s: one image row of u (%lld bytes) exceeds 4 GiB", what,
              (long long)W * Hd * 4);
  YAT_REQUIRE((int64_t)B * ytiles_of(H) < 65536, YAT_ERR_UNSUPPORTED, "%s: B * y-tiles = %lld exceeds the grid limit", what,
              (long long)B * ytiles_of(H));
  return YAT_OK;
}

// tensor map of a token-major [B, H, W, C] bf16 tensor with a {64 channels, 1 column, rows, 1} box
int image_tmap(CUtensorMap* out, const void* base, int B, int H, int W, int64_t C, int rows) {
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
  const uint32_t box[4] = {64u, 1u, (uint32_t)rows, 1u};
  return make_tmap_4d_bf16_plain(out, base, dims, box);
}
}  // namespace
}  // namespace yat

extern "C" int64_t yat_glumb_workspace_bytes(int B, int H, int Hd) {
  return (int64_t)B * yat::ytiles_of(H) * 2 * Hd * yat::kPart * (int64_t)sizeof(float);
}

extern "C" int yat_glumb_mid_fwd(const void* u, const void* w_dw, const void* b_dw, int B, int H, int W, int Hd,
                                 void* mid, void* c_out, void* stream) {
  using namespace yat;
  YAT_REQUIRE(u && w_dw && b_dw && mid, YAT_ERR_BAD_ARG, "yat_glumb_mid_fwd: null pointer argument");
  int rc = check_shape("yat_glumb_mid_fwd", B, H, W, Hd);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(u) && aligned16(mid) && aligned16(c_out), YAT_ERR_BAD_ARG,
              "yat_glumb_mid_fwd: tensors must be 16-byte aligned");
  GlumbArgs a{};
  if ((rc = image_tmap(&a.tm_u, u, B, H, W, 2 * (int64_t)Hd, kInCta)) != YAT_OK) return rc;
  a.p.w = (const __nv_bfloat16*)w_dw; a.p.bias = (const __nv_bfloat16*)b_dw;
  a.p.mid = (__nv_bfloat16*)mid; a.p.c = (__nv_bfloat16*)c_out; a.p.B = B; a.p.H = H; a.p.W = W; a.p.Hd = Hd;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(glumb_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem);
    attr_set = true;
  }
  glumb_fwd_kernel<<<dim3((Hd + 63) / 64, B * ytiles_of(H)), kThreads, kFwdSmem, (cudaStream_t)stream>>>(a);
  return check_launch("glumb_mid_fwd");
}

extern "C" int yat_glumb_mid_bwd(const void* u, const void* c, const void* w_dw, const void* dmid, int B, int H, int W,
                                 int Hd, void* du, void* dw, void* db, void* du_colsum, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(u && c && w_dw && dmid && du && dw && db && workspace, YAT_ERR_BAD_ARG,
              "yat_glumb_mid_bwd: null pointer argument");
  int rc = check_shape("yat_glumb_mid_bwd", B, H, W, Hd);
  if (rc != YAT_OK) return rc;
  YAT_REQUIRE(aligned16(u) && aligned16(c) && aligned16(dmid) && aligned16(du), YAT_ERR_BAD_ARG,
              "yat_glumb_mid_bwd: tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  GlumbArgs a{};
  if ((rc = image_tmap(&a.tm_u, u, B, H, W, 2 * (int64_t)Hd, kRowsCta)) != YAT_OK) return rc;
  if ((rc = image_tmap(&a.tm_c, c, B, H, W, 2 * (int64_t)Hd, kInCta)) != YAT_OK) return rc;
  if ((rc = image_tmap(&a.tm_d, dmid, B, H, W, (int64_t)Hd, kInCta)) != YAT_OK) return rc;
  a.p.w = (const __nv_bfloat16*)w_dw; a.p.du = (__nv_bfloat16*)du;
  a.p.partial = (float*)workspace; a.p.B = B; a.p.H = H; a.p.W = W; a.p.Hd = Hd;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(glumb_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem);
    attr_set = true;
  }
  const int parts = B * ytiles_of(H);
  glumb_bwd_kernel<<<dim3(2 * ((Hd + 63) / 64), parts), kThreads, kBwdSmem, st>>>(a);
  if ((rc = check_launch("glumb_bwd")) != YAT_OK) return rc;
  const int64_t n = (int64_t)2 * Hd * kPart;
  glumb_finish_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a.p.partial, parts, n, (__nv_bfloat16*)dw,
                                                                 (__nv_bfloat16*)db, (__nv_bfloat16*)du_colsum);
  return check_launch("glumb_finish");
}
