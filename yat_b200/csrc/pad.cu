// Ragged text embeddings -> padded batch + mask, on the device.
// Replaces the per-sample Python loop of F.pad / torch.zeros / stack and the pageable H2D copy of
// the padding zeros (train_sana.py:167-179, train_pixart_sigma.py:158-169): the host ships only the
// real tokens (packed [sum L_i, E]) and B+1 offsets.
#include "common.cuh"

namespace yat {
namespace {

__global__ void __launch_bounds__(256) pad_embeddings_kernel(const uint4* packed, const int32_t* offsets, int Lmax,
                                                             int vec_per_row, uint4* padded, int64_t* mask) {
  const int b = blockIdx.y;
  const int row = blockIdx.x;  // token slot in [0, Lmax)
  const int start = offsets[b];
  const int len = min(offsets[b + 1] - start, Lmax);
  const bool real = row < len;
  if (threadIdx.x == 0) mask[(int64_t)b * Lmax + row] = real ? 1 : 0;
  uint4* dst = padded + ((int64_t)b * Lmax + row) * vec_per_row;
  const uint4* src = packed + (int64_t)(start + row) * vec_per_row;
  for (int v = threadIdx.x; v < vec_per_row; v += blockDim.x)
    st_stream(dst + v, real ? ld_stream(src + v) : make_uint4(0, 0, 0, 0));
}

// ------------------------------------------------------------------------------------------------
// Head padding for the cross-attention SDPA (attention.py, _sdpa_key_bias_folded): the additive key-padding mask is
// folded into ONE extra head column, so the library attention runs without a mask tensor (B200, SANA-1.6B shape:
// cuDNN SDPA 207 -> 67 us forward, 757 -> 392 us forward + backward).
//   head_pad    x [rows, H*d]  ->  out [rows, H, da]:  out[.., :d] = x,  out[.., d] = tail (constant, or one value per
//               row: bias_row * sqrt(d) for the keys),  out[.., d+1:] = 0
//   head_unpad  xa [B, T, H, >= d] with element strides (sb, st, sh)  ->  out [B*T, H*d]   (slice + compaction)
// They are each other's backward.  One thread per 16-byte vector; pure streaming copies.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) head_pad_kernel(const uint4* __restrict__ x, const __nv_bfloat16* __restrict__ tail_row,
                                                       uint32_t tail_const, int64_t total_vec, int H, int dv, int dav,
                                                       uint4* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total_vec; i += (int64_t)gridDim.x * 256) {
    const int v = (int)(i % dav);
    const int64_t rh = i / dav;                      // row * H + h
    uint4 val = make_uint4(0u, 0u, 0u, 0u);
    if (v < dv) {
      val = ld_stream(x + rh * dv + v);
    } else if (v == dv) {
      val.x = tail_row ? (uint32_t)__bfloat16_as_ushort(tail_row[rh / H]) : tail_const;   // low half = column d
    }
    st_stream(out + i, val);
  }
}

__global__ void __launch_bounds__(256) head_unpad_kernel(const __nv_bfloat16* __restrict__ xa, int64_t sb, int64_t st,
                                                         int64_t sh, int T, int H, int dv, int64_t total_vec,
                                                         uint4* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total_vec; i += (int64_t)gridDim.x * 256) {
    const int v = (int)(i % dv);
    const int64_t rh = i / dv;
    const int h = (int)(rh % H);
    const int64_t row = rh / H;
    const int64_t b = row / T, t = row - b * T;
    st_stream(out + i, ld_stream(xa + b * sb + t * st + h * sh + (int64_t)v * 8));
  }
}

inline int copy_grid(int64_t total_vec) {
  const int64_t want = (total_vec + 255) / 256;
  return (int)(want < kNumSMs * 16 ? (want < 1 ? 1 : want) : kNumSMs * 16);
}

}  // namespace
}  // namespace yat

extern "C" int yat_head_pad(const void* x, const void* tail_row, float tail_const, int64_t rows, int H, int d, int da,
                            void* out, void* stream) {
  using namespace yat;
  YAT_REQUIRE(x && out, YAT_ERR_BAD_ARG, "yat_head_pad: null pointer argument");
  YAT_REQUIRE(rows > 0 && H > 0 && d > 0 && da > d, YAT_ERR_BAD_ARG, "yat_head_pad: rows=%lld H=%d d=%d da=%d",
              (long long)rows, H, d, da);
  YAT_REQUIRE(d % 8 == 0 && da % 8 == 0, YAT_ERR_UNSUPPORTED, "yat_head_pad: d=%d and da=%d must be multiples of 8", d, da);
  YAT_REQUIRE(((uintptr_t)x & 15u) == 0 && ((uintptr_t)out & 15u) == 0, YAT_ERR_BAD_ARG,
              "yat_head_pad: tensors must be 16-byte aligned");
  const int64_t total = rows * H * (da / 8);
  const uint32_t tc = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(tail_const));
  head_pad_kernel<<<copy_grid(total), 256, 0, (cudaStream_t)stream>>>((const uint4*)x, (const __nv_bfloat16*)tail_row, tc,
                                                                     total, H, d / 8, da / 8, (uint4*)out);
  return check_launch("head_pad");
}

extern "C" int yat_head_unpad(const void* xa, int64_t stride_b, int64_t stride_t, int64_t stride_h, int B, int T, int H,
                              int d, void* out, void* stream) {
  using namespace yat;
  YAT_REQUIRE(xa && out, YAT_ERR_BAD_ARG, "yat_head_unpad: null pointer argument");
  YAT_REQUIRE(B > 0 && T > 0 && H > 0 && d > 0, YAT_ERR_BAD_ARG, "yat_head_unpad: B=%d T=%d H=%d d=%d", B, T, H, d);
  YAT_REQUIRE(d % 8 == 0 && stride_b % 8 == 0 && stride_t % 8 == 0 && stride_h % 8 == 0, YAT_ERR_UNSUPPORTED,
              "yat_head_unpad: d and the strides must be multiples of 8 elements");
  YAT_REQUIRE(((uintptr_t)xa & 15u) == 0 && ((uintptr_t)out & 15u) == 0, YAT_ERR_BAD_ARG,
              "yat_head_unpad: tensors must be 16-byte aligned");
  const int64_t total = (int64_t)B * T * H * (d / 8);
  head_unpad_kernel<<<copy_grid(total), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)xa, stride_b, stride_t,
                                                                       stride_h, T, H, d / 8, total, (uint4*)out);
  return check_launch("head_unpad");
}

extern "C" int yat_pad_embeddings(const void* packed, const int32_t* offsets, int B, int Lmax, int E, void* padded,
                                  int64_t* mask, void* stream) {
  using namespace yat;
  YAT_REQUIRE(packed && offsets && padded && mask, YAT_ERR_BAD_ARG, "yat_pad_embeddings: null pointer argument");
  YAT_REQUIRE(B > 0 && Lmax > 0 && E > 0 && B <= 65535, YAT_ERR_BAD_ARG, "yat_pad_embeddings: bad dimensions");
  YAT_REQUIRE(E % 8 == 0, YAT_ERR_UNSUPPORTED, "yat_pad_embeddings: E=%d must be a multiple of 8", E);
  YAT_REQUIRE(((uintptr_t)packed & 15u) == 0 && ((uintptr_t)padded & 15u) == 0, YAT_ERR_BAD_ARG,
              "yat_pad_embeddings: tensors must be 16-byte aligned");
  pad_embeddings_kernel<<<dim3(Lmax, B), 256, 0, (cudaStream_t)stream>>>((const uint4*)packed, offsets, Lmax, E / 8,
                                                                       (uint4*)padded, mask);
  return check_launch("pad_embeddings");
}
