// (b) ResNet stem (7x7 stride-2 pad-3 convolution on 3 input channels) for the tensor-core path.
//
// With C=3 a channel run is 6 bytes - below TMA's 16-byte granule - so the stem cannot use the
// shifted-box trick of conv_tc.cu. Instead one HBM-bound kernel writes the im2col matrix
//   col[pixel][(r*7+s)*3+c]  (147 real columns, zero-padded to 192 = 3 x 64 bf16 = three 128-byte rows)
// straight from the caller's image tensor (any strides, fp32 or bf16: the NCHW->NHWC conversion and the
// bf16 cast are folded in), after which the stem is a 1x1 convolution with 192 input channels and runs
// on the same tcgen05 kernels (fprop and wgrad).
#include "common.cuh"

namespace iit {

constexpr int STEM_K = 192;

// One block per (image, output row): the 7 input rows x 3 channels that the row needs are staged in shared
// memory with coalesced row reads (zero halo included), then every thread emits 16-byte groups of 8 im2col
// columns; consecutive threads write consecutive groups, i.e. whole 384-byte im2col rows back to back.
template <typename T>
__global__ void __launch_bounds__(256) stem_im2col_kernel(const T* __restrict__ x, int64_t sn, int64_t sc, int64_t sh, int64_t sw,
                                                          __nv_bfloat16* __restrict__ col, int N, int H, int W, int P, int Q) {
  extern __shared__ float tile[];           // [3][7][W + 6]
  const int WP = W + 6;
  const int n = blockIdx.x / P, p = blockIdx.x - n * P;
  for (int i = threadIdx.x; i < 21 * WP; i += blockDim.x) {
    const int cr = i / WP, wp = i - cr * WP;
    const int c = cr / 7, r = cr - c * 7;
    const int ih = p * 2 - 3 + r, iw = wp - 3;
    float v = 0.f;
    if (ih >= 0 && ih < H && iw >= 0 && iw < W) v = to_f<T>(x[n * sn + c * sc + ih * sh + iw * sw]);
    tile[i] = v;
  }
  __syncthreads();
  __nv_bfloat16* out = col + (int64_t)blockIdx.x * Q * STEM_K;
  for (int i = threadIdx.x; i < Q * (STEM_K / 8); i += blockDim.x) {
    const int q = i / (STEM_K / 8), g = i - q * (STEM_K / 8);
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int kk = g * 8 + j;
      float val = 0.f;
      if (kk < 147) {
        const int r = kk / 21, rem = kk - r * 21;
        const int s = rem / 3, c = rem - s * 3;
        val = tile[(c * 7 + r) * WP + q * 2 + s];
      }
      v[j] = val;
    }
    __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]), h1 = __floats2bfloat162_rn(v[2], v[3]);
    __nv_bfloat162 h2 = __floats2bfloat162_rn(v[4], v[5]), h3 = __floats2bfloat162_rn(v[6], v[7]);
    uint4 o;
    o.x = *reinterpret_cast<uint32_t*>(&h0); o.y = *reinterpret_cast<uint32_t*>(&h1);
    o.z = *reinterpret_cast<uint32_t*>(&h2); o.w = *reinterpret_cast<uint32_t*>(&h3);
    *reinterpret_cast<uint4*>(out + (int64_t)i * 8) = o;
  }
}

}  // namespace iit

using namespace iit;

extern "C" int iit_stem_im2col(const void* x, int dtype, const int64_t* strides_nchw, void* col, int n, int h, int w,
                               void* stream) {
  IIT_REQUIRE(x && col && strides_nchw, "iit_stem_im2col: null pointer");
  IIT_REQUIRE((reinterpret_cast<uintptr_t>(col) & 15) == 0, "iit_stem_im2col: col must be 16-byte aligned");
  const int P = (h + 6 - 7) / 2 + 1, Q = (w + 6 - 7) / 2 + 1;
  const size_t smem = (size_t)21 * (w + 6) * sizeof(float);
  IIT_REQUIRE(smem <= 48 * 1024, "iit_stem_im2col: image width %d too large for the staging tile", w);
  IIT_REQUIRE((int64_t)n * P < (1ll << 31), "iit_stem_im2col: too many output rows");
  cudaStream_t st = as_stream(stream);
  IIT_DISPATCH_DTYPE(dtype, T, {
    stem_im2col_kernel<T><<<(unsigned)(n * P), 256, smem, st>>>((const T*)x, strides_nchw[0], strides_nchw[1], strides_nchw[2],
                                                                strides_nchw[3], (__nv_bfloat16*)col, n, h, w, P, Q);
  })
  IIT_LAUNCH_CHECK("iit_stem_im2col");
  return 0;
}
