// REPA cosine alignment, forward + backward.
//   z_pool[b,t,:] = mean over the adaptive_avg_pool2d window of teacher cell t of z[b,(r,c),:]
//   L = -(1/(B*T)) sum_{b,t} cos(z_pool, y)          cos uses F.cosine_similarity's eps clamp
//   dz[b,(r,c),:] = gscale * sum_{t : (r,c) in window(t)} dL/dz_pool[b,t,:] / |window(t)|
// Specified in SURVEY.md §8 A8 (the reference's calculate_loss, common/repa.py:137-154, is a
// no-grad placeholder).  One warp per teacher token; when the windows tile the student grid
// exactly (gh % th == 0 and gw % tw == 0, e.g. 32x32 -> 16x16) the gradient is written in the
// same pass; otherwise (22x44 and the other aspect buckets: overlapping windows) the pooled
// gradient is staged in fp32 and a second kernel gathers it per student token.
#include "common.cuh"

namespace yat {
namespace {

constexpr int kWarpsPerBlock = 8;
constexpr int kThreads = kWarpsPerBlock * 32;
constexpr int kMaxBlocks = 4096;
constexpr float kCosEps = 1e-8f;

struct RepaParams {
  const __nv_bfloat16* z;
  const void* y;
  int y_is_f32;
  int B, gh, gw, th, tw, E;
  float gscale;
  float* loss_out;
  __nv_bfloat16* dz;
  float* gbuf;        // [B, th*tw, E] fp32, overlapping-window path only
  float* partials;    // [gridDim.x]
  unsigned int* counter;
  int direct;         // windows tile the grid exactly -> write dz in phase A
  int total_warps;
};

__device__ __forceinline__ void window(int i, int n_in, int n_out, int& s, int& e) {
  s = (i * n_in) / n_out;
  e = ((i + 1) * n_in + n_out - 1) / n_out;
}

__device__ __forceinline__ void unpack8(const uint4& v, float f[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}

// VPL = 128-bit vectors of 8 channels per lane: E <= VPL*256.  QUAD: every window is exactly 2x2 student tokens
// (32x32 -> 16x16, the SANA 1024 px case): the four rows are fetched up front, 4 * VPL loads in flight per lane.
template <int VPL, bool QUAD>
__global__ void __launch_bounds__(kThreads) repa_pool_cos_kernel(const RepaParams p) {
  __shared__ float s_warp[kWarpsPerBlock];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int T = p.th * p.tw;
  const int nvec = p.E >> 3;
  float cos_acc = 0.f;

  for (int item = blockIdx.x * kWarpsPerBlock + warp; item < p.total_warps; item += gridDim.x * kWarpsPerBlock) {
    const int b = item / T, t = item - b * T;
    const int ti = t / p.tw, tj = t - ti * p.tw;
    int r0, r1, c0, c1;
    window(ti, p.gh, p.th, r0, r1);
    window(tj, p.gw, p.tw, c0, c1);
    const int cnt = (r1 - r0) * (c1 - c0);
    const float inv_cnt = 1.0f / (float)cnt;

    float acc[VPL][8];
#pragma unroll
    for (int j = 0; j < VPL; ++j)
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[j][k] = 0.f;
    if (QUAD) {
      uint4 q[4][VPL];
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const uint4* row = reinterpret_cast<const uint4*>(
            p.z + ((int64_t)b * p.gh * p.gw + (int64_t)(r0 + (w >> 1)) * p.gw + c0 + (w & 1)) * p.E);
#pragma unroll
        for (int j = 0; j < VPL; ++j)
          if (lane + 32 * j < nvec) q[w][j] = ld_stream(row + lane + 32 * j);
      }
#pragma unroll
      for (int w = 0; w < 4; ++w)                                  // same summation order as the generic loop
#pragma unroll
        for (int j = 0; j < VPL; ++j)
          if (lane + 32 * j < nvec) {
            float f[8];
            unpack8(q[w][j], f);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[j][k] += f[k];
          }
    } else {
      for (int r = r0; r < r1; ++r)
        for (int c = c0; c < c1; ++c) {
          const uint4* row = reinterpret_cast<const uint4*>(p.z + ((int64_t)b * p.gh * p.gw + (int64_t)r * p.gw + c) * p.E);
#pragma unroll
          for (int j = 0; j < VPL; ++j) {
            const int v = lane + 32 * j;
            if (v < nvec) {
              float f[8];
              unpack8(ld_stream(row + v), f);
#pragma unroll
              for (int k = 0; k < 8; ++k) acc[j][k] += f[k];
            }
          }
        }
    }
    float yv[VPL][8];
    float dot = 0.f, nz = 0.f, ny = 0.f;
    const int64_t yoff = ((int64_t)b * T + t) * p.E;
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
      const int v = lane + 32 * j;
      if (v < nvec) {
        if (p.y_is_f32) {
          const float4* yr = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.y) + yoff) + 2 * v;
          const float4 a = __ldg(yr), c = __ldg(yr + 1);
          yv[j][0] = a.x; yv[j][1] = a.y; yv[j][2] = a.z; yv[j][3] = a.w;
          yv[j][4] = c.x; yv[j][5] = c.y; yv[j][6] = c.z; yv[j][7] = c.w;
        } else {
          unpack8(ld_stream(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.y) + yoff) + v), yv[j]);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          acc[j][k] *= inv_cnt;
          dot = fmaf(acc[j][k], yv[j][k], dot);
          nz = fmaf(acc[j][k], acc[j][k], nz);
          ny = fmaf(yv[j][k], yv[j][k], ny);
        }
      }
    }
    dot = warp_sum(dot);
    nz = warp_sum(nz);
    ny = warp_sum(ny);
    const float rz = sqrtf(nz), ry = sqrtf(ny);
    const float nzc = fmaxf(rz, kCosEps), nyc = fmaxf(ry, kCosEps);
    const float cosv = dot / (nzc * nyc);
    if (lane == 0) cos_acc += cosv;

    if (p.dz) {
      // dL/dz_pool = -(gscale/(B*T)) * ( y/(|z||y|) - cos * z_pool/|z|^2 ), spread over the window
      const float s = -p.gscale / ((float)p.B * (float)T) * inv_cnt;
      const float a = s / (nzc * nyc);
      const float bcoef = rz > kCosEps ? s * cosv / (nzc * rz) : 0.f;
      if (p.direct) {
        for (int r = r0; r < r1; ++r)
          for (int c = c0; c < c1; ++c) {
            uint4* row = reinterpret_cast<uint4*>(p.dz + ((int64_t)b * p.gh * p.gw + (int64_t)r * p.gw + c) * p.E);
#pragma unroll
            for (int j = 0; j < VPL; ++j) {
              const int v = lane + 32 * j;
              if (v < nvec) {
                float g[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) g[k] = a * yv[j][k] - bcoef * acc[j][k];
                st_stream(row + v, make_uint4(pack_bf16(g[0], g[1]), pack_bf16(g[2], g[3]), pack_bf16(g[4], g[5]),
                                              pack_bf16(g[6], g[7])));
              }
            }
          }
      } else {
        float4* grow = reinterpret_cast<float4*>(p.gbuf + yoff);
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
          const int v = lane + 32 * j;
          if (v < nvec) {
            float g[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] = a * yv[j][k] - bcoef * acc[j][k];
            grow[2 * v] = make_float4(g[0], g[1], g[2], g[3]);
            grow[2 * v + 1] = make_float4(g[4], g[5], g[6], g[7]);
          }
        }
      }
    }
  }

  // deterministic two-stage reduction of sum(cos)
  if (lane == 0) s_warp[warp] = cos_acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float bsum = 0.f;
    for (int w = 0; w < kWarpsPerBlock; ++w) bsum += s_warp[w];
    p.partials[blockIdx.x] = bsum;
    __threadfence();
    s_last = atomicAdd(p.counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (warp == 0) {
    float s = 0.f;
    for (int i = lane; i < (int)gridDim.x; i += 32) s += __ldcg(p.partials + i);
    s = warp_sum(s);
    if (lane == 0) {
      p.loss_out[0] = -s / ((float)p.B * (float)T);
      *p.counter = 0u;
    }
  }
}

// overlapping windows: dz[b,(r,c),:] = sum over covering teacher cells of the staged gradient
__global__ void __launch_bounds__(kThreads) repa_gather_kernel(const RepaParams p) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n_tok = p.B * p.gh * p.gw;
  const int nvec = p.E >> 3;
  const int T = p.th * p.tw;
  for (int tok = blockIdx.x * kWarpsPerBlock + warp; tok < n_tok; tok += gridDim.x * kWarpsPerBlock) {
    const int b = tok / (p.gh * p.gw);
    const int rc = tok - b * p.gh * p.gw;
    const int r = rc / p.gw, c = rc - r * p.gw;
    uint4* out = reinterpret_cast<uint4*>(p.dz + (int64_t)tok * p.E);
    for (int v = lane; v < nvec; v += 32) {
      float g[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int ti = 0; ti < p.th; ++ti) {
        int r0, r1;
        window(ti, p.gh, p.th, r0, r1);
        if (r < r0 || r >= r1) continue;
        for (int tj = 0; tj < p.tw; ++tj) {
          int c0, c1;
          window(tj, p.gw, p.tw, c0, c1);
          if (c < c0 || c >= c1) continue;
          const float4* src = reinterpret_cast<const float4*>(p.gbuf + ((int64_t)b * T + ti * p.tw + tj) * p.E) + 2 * v;
          const float4 a = __ldcg(src), d = __ldcg(src + 1);
          g[0] += a.x; g[1] += a.y; g[2] += a.z; g[3] += a.w;
          g[4] += d.x; g[5] += d.y; g[6] += d.z; g[7] += d.w;
        }
      }
      st_stream(out + v, make_uint4(pack_bf16(g[0], g[1]), pack_bf16(g[2], g[3]), pack_bf16(g[4], g[5]),
                                    pack_bf16(g[6], g[7])));
    }
  }
}

inline int64_t repa_header_bytes() { return (int64_t)(kMaxBlocks + 4) * sizeof(float); }

}  // namespace
}  // namespace yat

extern "C" int64_t yat_repa_workspace_bytes(int B, int th, int tw, int E) {
  return yat::repa_header_bytes() + (int64_t)B * th * tw * E * (int64_t)sizeof(float);
}

extern "C" int yat_repa_cos(const void* z, const void* y, int y_dtype, int B, int gh, int gw, int th, int tw, int E,
                            float gscale, float* loss_out, void* dz, void* workspace, void* stream) {
  using namespace yat;
  YAT_REQUIRE(z && y && loss_out && workspace, YAT_ERR_BAD_ARG, "yat_repa_cos: null pointer argument");
  YAT_REQUIRE(B > 0 && gh > 0 && gw > 0 && th > 0 && tw > 0 && E > 0, YAT_ERR_BAD_ARG,
              "yat_repa_cos: non-positive dimension");
  YAT_REQUIRE(E % 8 == 0 && E <= 2048, YAT_ERR_UNSUPPORTED, "yat_repa_cos: E=%d must be a multiple of 8 and <= 2048", E);
  YAT_REQUIRE(y_dtype == YAT_DTYPE_BF16 || y_dtype == YAT_DTYPE_F32, YAT_ERR_UNSUPPORTED,
              "yat_repa_cos: teacher dtype %d not supported", y_dtype);
  YAT_REQUIRE(((uintptr_t)z & 15u) == 0 && ((uintptr_t)y & 15u) == 0 && ((uintptr_t)dz & 15u) == 0, YAT_ERR_BAD_ARG,
              "yat_repa_cos: tensors must be 16-byte aligned");
  RepaParams p{};
  p.z = (const __nv_bfloat16*)z; p.y = y; p.y_is_f32 = y_dtype == YAT_DTYPE_F32;
  p.B = B; p.gh = gh; p.gw = gw; p.th = th; p.tw = tw; p.E = E; p.gscale = gscale;
  p.loss_out = loss_out; p.dz = (__nv_bfloat16*)dz;
  p.partials = (float*)workspace;
  p.counter = (unsigned int*)((float*)workspace + kMaxBlocks);
  p.gbuf = (float*)((char*)workspace + repa_header_bytes());
  p.direct = (gh % th == 0) && (gw % tw == 0);
  p.total_warps = B * th * tw;
  const int need = (p.total_warps + kWarpsPerBlock - 1) / kWarpsPerBlock;
  const int grid = need < kMaxBlocks ? need : kMaxBlocks;
  cudaStream_t st = (cudaStream_t)stream;
  const bool quad = gh == 2 * th && gw == 2 * tw;
  if (E <= 768 && quad)  repa_pool_cos_kernel<3, true><<<grid, kThreads, 0, st>>>(p);
  else if (E <= 768)     repa_pool_cos_kernel<3, false><<<grid, kThreads, 0, st>>>(p);
  else if (E <= 1024)    repa_pool_cos_kernel<4, false><<<grid, kThreads, 0, st>>>(p);
  else                   repa_pool_cos_kernel<8, false><<<grid, kThreads, 0, st>>>(p);
  int rc = check_launch("repa_pool_cos");
  if (rc != YAT_OK) return rc;
  if (dz && !p.direct) {
    const int n_tok = B * gh * gw;
    const int g2 = (n_tok + kWarpsPerBlock - 1) / kWarpsPerBlock;
    repa_gather_kernel<<<g2 < 65535 ? g2 : 65535, kThreads, 0, st>>>(p);
    rc = check_launch("repa_gather");
  }
  return rc;
}
