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

// ---- space-to-depth form of the stem (conv_win.cu runs the convolution) ---------------------------------------
// One block per (image, s2d row): the two padded input rows x 3 channels are staged in shared memory with
// coalesced reads, then each thread writes one s2d pixel's 16 channels (two 16-byte stores; 32-byte rows back to back).
template <typename T>
__global__ void __launch_bounds__(128) stem_s2d_kernel(const T* __restrict__ x, int64_t sn, int64_t sc, int64_t sh, int64_t sw,
                                                       __nv_bfloat16* __restrict__ out, int H, int W, int SH, int SW) {
  extern __shared__ float tile[];            // [3 channels][2 rows][W + 6]
  const int WP = W + 6;
  const int n = blockIdx.x / SH, sy = blockIdx.x - n * SH;
  for (int i = threadIdx.x; i < 6 * WP; i += blockDim.x) {
    const int ca = i / WP, xp = i - ca * WP;
    const int c = ca >> 1, a = ca & 1;
    const int ih = 2 * sy + a - 3, iw = xp - 3;
    float v = 0.f;
    if (ih >= 0 && ih < H && iw >= 0 && iw < W) v = to_f<T>(x[n * sn + c * sc + ih * sh + iw * sw]);
    tile[i] = v;
  }
  __syncthreads();
  __nv_bfloat16* row = out + (int64_t)blockIdx.x * SW * 16;
  for (int sx = threadIdx.x; sx < SW; sx += blockDim.x) {
    float v[16];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int c = 0; c < 3; ++c) v[(a * 2 + b) * 3 + c] = tile[(c * 2 + a) * WP + 2 * sx + b];
    v[12] = v[13] = v[14] = v[15] = 0.f;
    uint4 o[2];
    uint32_t* w32 = reinterpret_cast<uint32_t*>(o);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      __nv_bfloat162 h2 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
      w32[j] = *reinterpret_cast<uint32_t*>(&h2);
    }
    uint4* dst = reinterpret_cast<uint4*>(row + (int64_t)sx * 16);
    dst[0] = o[0]; dst[1] = o[1];
  }
}

// Whole image per block (images up to 63 x 63): the image is staged in shared memory with flat coalesced reads (16-byte
// loads when it is dense NCHW fp32), then the s2d pixels - 32 contiguous bytes each, the whole output image one contiguous
// run - are written back to back.
template <typename T>
__global__ void __launch_bounds__(256) stem_s2d_image_kernel(const T* __restrict__ x, int64_t sn, int64_t sc, int64_t sh, int64_t sw,
                                                             __nv_bfloat16* __restrict__ out, int N, int H, int W, int SH, int SW,
                                                             int dense) {
  extern __shared__ float img[];             // [3][H][W]
  const int HW = H * W;
  for (int n = blockIdx.x; n < N; n += gridDim.x) {
    const T* src = x + (int64_t)n * sn;
    if (dense && sizeof(T) == 4 && (HW * 3) % 4 == 0) {
      const float4* s4 = reinterpret_cast<const float4*>(src);
      for (int i = threadIdx.x; i < HW * 3 / 4; i += blockDim.x) reinterpret_cast<float4*>(img)[i] = s4[i];
    } else {
      for (int i = threadIdx.x; i < HW * 3; i += blockDim.x) {
        const int c = i / HW, r = i - c * HW, h = r / W, w = r - h * W;
        img[i] = to_f<T>(src[c * sc + h * sh + w * sw]);
      }
    }
    __syncthreads();
    uint4* dst = reinterpret_cast<uint4*>(out + (int64_t)n * SH * SW * 16);
    for (int i = threadIdx.x; i < SH * SW * 2; i += blockDim.x) {     // two 16-byte halves per s2d pixel
      const int pix = i >> 1, half = i & 1;
      const int sy = pix / SW, sx = pix - sy * SW;
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int ch = half * 8 + j;         // (a*2+b)*3+c for ch < 12, zero above
        float val = 0.f;
        if (ch < 12) {
          const int ab = ch / 3, c = ch - ab * 3, a = ab >> 1, b = ab & 1;
          const int ih = 2 * sy + a - 3, iw = 2 * sx + b - 3;
          if (ih >= 0 && ih < H && iw >= 0 && iw < W) val = img[c * HW + ih * W + iw];
        }
        v[j] = val;
      }
      uint4 o;
      uint32_t* w32 = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        __nv_bfloat162 h2 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
        w32[j] = *reinterpret_cast<uint32_t*>(&h2);
      }
      dst[i] = o;
    }
    __syncthreads();
  }
}

// w [64,3,7,7] -> w2 [64][R 4][S 4][16]: w2[k,R,S,(a*2+b)*3+c] = w[k,c,2R+a,2S+b] (zero where 2R+a or 2S+b is 7)
__global__ void stem_filter_s2d_kernel(const float* __restrict__ w, int64_t sk, int64_t sc, int64_t sr, int64_t ss,
                                       __nv_bfloat16* __restrict__ w2) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * 256) return;
  const int ch = i & 15, S = (i >> 4) & 3, R = (i >> 6) & 3, k = i >> 8;
  float v = 0.f;
  if (ch < 12) {
    const int ab = ch / 3, c = ch - ab * 3, a = ab >> 1, b = ab & 1;
    const int r = 2 * R + a, s = 2 * S + b;
    if (r < 7 && s < 7) v = w[k * sk + c * sc + r * sr + s * ss];
  }
  w2[i] = __float2bfloat16_rn(v);
}

}  // namespace iit

using namespace iit;

extern "C" int iit_stem_s2d(const void* x, int dtype, const int64_t* strides_nchw, void* s2d, int n, int h, int w, void* stream) {
  IIT_REQUIRE(x && s2d && strides_nchw, "iit_stem_s2d: null pointer");
  IIT_REQUIRE(h % 2 == 0 && w % 2 == 0 && h >= 2 && w >= 2, "iit_stem_s2d: image height and width must be even (got %dx%d)", h, w);
  IIT_REQUIRE((reinterpret_cast<uintptr_t>(s2d) & 15) == 0, "iit_stem_s2d: output must be 16-byte aligned");
  const int SH = (h + 6) / 2, SW = (w + 6) / 2;
  const size_t smem = (size_t)6 * (w + 6) * sizeof(float);
  IIT_REQUIRE(smem <= 48 * 1024, "iit_stem_s2d: image width %d too large for the staging tile", w);
  IIT_REQUIRE((int64_t)n * SH < (1ll << 31), "iit_stem_s2d: too many rows");
  cudaStream_t st = as_stream(stream);
  const size_t img_smem = (size_t)3 * h * w * sizeof(float);
  if (img_smem <= 48 * 1024) {
    const int dense = strides_nchw[3] == 1 && strides_nchw[2] == w && strides_nchw[1] == (int64_t)h * w &&
                      (reinterpret_cast<uintptr_t>(x) & 15) == 0 && ((strides_nchw[0] * (dtype == IIT_F32 ? 4 : 2)) & 15) == 0;
    const int grid = n < kNumSMs * 4 ? n : kNumSMs * 4;
    IIT_DISPATCH_DTYPE(dtype, T, {
      stem_s2d_image_kernel<T><<<grid, 256, img_smem, st>>>((const T*)x, strides_nchw[0], strides_nchw[1], strides_nchw[2],
                                                           strides_nchw[3], (__nv_bfloat16*)s2d, n, h, w, SH, SW, dense);
    })
    IIT_LAUNCH_CHECK("iit_stem_s2d");
    return 0;
  }
  IIT_DISPATCH_DTYPE(dtype, T, {
    stem_s2d_kernel<T><<<(unsigned)(n * SH), 128, smem, st>>>((const T*)x, strides_nchw[0], strides_nchw[1], strides_nchw[2],
                                                             strides_nchw[3], (__nv_bfloat16*)s2d, h, w, SH, SW);
  })
  IIT_LAUNCH_CHECK("iit_stem_s2d");
  return 0;
}

extern "C" int iit_stem_filter_s2d(const float* w, const int64_t* strides_kcrs, void* w2, void* stream) {
  IIT_REQUIRE(w && strides_kcrs && w2, "iit_stem_filter_s2d: null pointer");
  stem_filter_s2d_kernel<<<64, 256, 0, as_stream(stream)>>>(w, strides_kcrs[0], strides_kcrs[1], strides_kcrs[2], strides_kcrs[3],
                                                           (__nv_bfloat16*)w2);
  IIT_LAUNCH_CHECK("iit_stem_filter_s2d");
  return 0;
}

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
