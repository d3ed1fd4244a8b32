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

}  // namespace
}  // namespace yat

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
