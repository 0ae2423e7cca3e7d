// (b) ResNet-18 convolutions as NHWC implicit GEMMs on the CUDA cores (fp32 FFMA accumulate).
//
// This is the exact-fp32 path (parity within 1e-3 of the fp32 reference needs more mantissa than
// TF32/bf16 tensor-core inputs give) and the catch-all for shapes the tcgen05 path does not take
// (7x7 stem with C=3, strided and 1x1 convolutions, weight gradients).
//
//   fprop : M = n*p*q pixels, N = k filters,  K = r*s*c      A = im2col(x) gathered, B = w[k][rsc]
//   dgrad : M = n*h*w pixels, N = c channels, K = r*s*k      A = col2im gather of dy, B = w_t[c][rsk]
//   wgrad : M = r*s*c,        N = k filters,  K = n*p*q      A = im2col(x)^T gathered, B = dy
//
// Tile 128(M) x 64(N) x 16(K), 256 threads, 8x4 register micro-tile, register-staged double
// buffering (next tile's global loads are in flight during the FMAs). Channels are innermost in
// NHWC, so when the channel count is a multiple of 16 every 16-wide K slice of A is one contiguous
// 64 B run per pixel and is fetched with 16 B vector loads; otherwise (the 3-channel stem) a scalar
// gather with full index decode is used.
#include <atomic>

#include "common.cuh"

namespace iit {

constexpr int BM = 128, BN = 64, BK = 16;
constexpr int APAD = 4, BPAD = 4;

struct ConvP {
  int N, Hi, Wi, Ci;   // gathered tensor (x for fprop/wgrad, dy for dgrad)
  int Ho, Wo, Co;      // produced pixel grid / channels (y for fprop, dx for dgrad)
  int R, S, stride, pad;
  int64_t M;           // rows of the GEMM
  int Kg;              // reduction length
};

// one 16-wide K slice of the 8x4 micro-tile update
__device__ __forceinline__ void mma_tile(const float (*As)[BM + APAD], const float (*Bs)[BN + BPAD], int ty, int tx,
                                         float (&acc)[8][4]) {
#pragma unroll
  for (int kk = 0; kk < BK; ++kk) {
    float a[8], b[4];
    float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 8]);
    float4 a1 = *reinterpret_cast<const float4*>(&As[kk][ty * 8 + 4]);
    float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
    a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
    b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w;
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
  }
}

// ------------------------------------------------------------------------------------------------
// fprop / dgrad
// ------------------------------------------------------------------------------------------------
template <typename T, bool DGRAD, bool VEC>
__global__ void __launch_bounds__(256) conv_gemm_kernel(const T* __restrict__ in, const T* __restrict__ wt,
                                                        T* __restrict__ out, ConvP p, double* __restrict__ bn_sum,
                                                        double* __restrict__ bn_sq) {
  __shared__ __align__(16) float As[BK][BM + APAD];
  __shared__ __align__(16) float Bs[BK][BN + BPAD];
  __shared__ float s_red[2][BN];
  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;

  // A loader: row = t>>1, k offsets [half*8, half*8+8)
  const int arow = t >> 1, ahalf = t & 1;
  const int64_t am = m0 + arow;
  const bool arow_ok = am < p.M;
  int an = 0, aoh = 0, aow = 0;
  if (arow_ok) {
    int64_t hw = (int64_t)p.Ho * p.Wo;
    an = (int)(am / hw);
    int rem = (int)(am - (int64_t)an * hw);
    aoh = rem / p.Wo;
    aow = rem - aoh * p.Wo;
  }
  // B loader: row = t>>2 (filter / channel), k offsets [kq*4, kq*4+4)
  const int brow = t >> 2, bkq = t & 3;
  const int bco = n0 + brow;
  const bool brow_ok = bco < p.Co;

  float areg[8], breg[4];
  auto tap = [&](int r, int s, int& ih, int& iw) -> bool {
    if (!DGRAD) {
      ih = aoh * p.stride - p.pad + r;
      iw = aow * p.stride - p.pad + s;
      return ih >= 0 && ih < p.Hi && iw >= 0 && iw < p.Wi;
    } else {
      int th = aoh + p.pad - r, tw = aow + p.pad - s;
      if (th < 0 || tw < 0) return false;
      ih = th / p.stride; iw = tw / p.stride;
      return (th - ih * p.stride == 0) && (tw - iw * p.stride == 0) && ih < p.Hi && iw < p.Wi;
    }
  };
  auto load_tiles = [&](int k0) {
    if (VEC) {
      int rs = k0 / p.Ci, c0 = k0 - rs * p.Ci;
      int r = rs / p.S, s = rs - r * p.S;
      int ih, iw;
      bool ok = arow_ok && tap(r, s, ih, iw);
      if (ok) {
        const T* src = in + (((int64_t)an * p.Hi + ih) * p.Wi + iw) * p.Ci + c0 + ahalf * 8;
        float v0[4], v1[4];
        Vec4<T>::load(src, v0);
        Vec4<T>::load(src + 4, v1);
#pragma unroll
        for (int j = 0; j < 4; ++j) { areg[j] = v0[j]; areg[4 + j] = v1[j]; }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) areg[j] = 0.f;
      }
      if (brow_ok) {
        float v[4];
        Vec4<T>::load(wt + (int64_t)bco * p.Kg + k0 + bkq * 4, v);
#pragma unroll
        for (int j = 0; j < 4; ++j) breg[j] = v[j];
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) breg[j] = 0.f;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        int k = k0 + ahalf * 8 + j;
        float v = 0.f;
        if (arow_ok && k < p.Kg) {
          int rs = k / p.Ci, c = k - rs * p.Ci;
          int r = rs / p.S, s = rs - r * p.S;
          int ih, iw;
          if (tap(r, s, ih, iw)) v = to_f<T>(in[(((int64_t)an * p.Hi + ih) * p.Wi + iw) * p.Ci + c]);
        }
        areg[j] = v;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        int k = k0 + bkq * 4 + j;
        breg[j] = (brow_ok && k < p.Kg) ? to_f<T>(wt[(int64_t)bco * p.Kg + k]) : 0.f;
      }
    }
  };
  auto store_tiles = [&]() {
#pragma unroll
    for (int j = 0; j < 8; ++j) As[ahalf * 8 + j][arow] = areg[j];
#pragma unroll
    for (int j = 0; j < 4; ++j) Bs[bkq * 4 + j][brow] = breg[j];
  };

  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int nk = (p.Kg + BK - 1) / BK;
  load_tiles(0);
  store_tiles();
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    if (kt + 1 < nk) load_tiles((kt + 1) * BK);
    mma_tile(As, Bs, ty, tx, acc);
    __syncthreads();
    if (kt + 1 < nk) {
      store_tiles();
      __syncthreads();
    }
  }

  // epilogue: rows m0+ty*8+i, cols n0+tx*4+j
  const int co = n0 + tx * 4;
  float csum[4] = {0.f, 0.f, 0.f, 0.f}, csq[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int64_t m = m0 + ty * 8 + i;
    if (m < p.M) {
      if (co + 3 < p.Co && (p.Co & 3) == 0) {
        float v[4] = {acc[i][0], acc[i][1], acc[i][2], acc[i][3]};
        Vec4<T>::store(out + m * p.Co + co, v);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (co + j < p.Co) out[m * p.Co + co + j] = from_f<T>(acc[i][j]);
      }
      if (bn_sum) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float v = to_f<T>(from_f<T>(acc[i][j]));  // statistics of the values as stored
          csum[j] += v; csq[j] += v * v;
        }
      }
    }
  }
  if (bn_sum) {  // block-uniform branch
    if (t < BN) { s_red[0][t] = 0.f; s_red[1][t] = 0.f; }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      // lanes l and l^16 hold the same columns (ty differs by one)
      float a = csum[j] + __shfl_xor_sync(0xffffffffu, csum[j], 16);
      float b = csq[j] + __shfl_xor_sync(0xffffffffu, csq[j], 16);
      if ((t & 16) == 0) { atomicAdd(&s_red[0][tx * 4 + j], a); atomicAdd(&s_red[1][tx * 4 + j], b); }
    }
    __syncthreads();
    if (t < BN && n0 + t < p.Co) {
      atomicAdd(&bn_sum[n0 + t], (double)s_red[0][t]);
      atomicAdd(&bn_sq[n0 + t], (double)s_red[1][t]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// wgrad: out[z][co][m] partial over the pixel range of split z   (m = (r*S+s)*C + c)
// ------------------------------------------------------------------------------------------------
template <typename T, bool VEC>
__global__ void __launch_bounds__(256) wgrad_kernel(const T* __restrict__ x, const T* __restrict__ dy,
                                                    float* __restrict__ out, ConvP p, int64_t pix_per_split) {
  // here p.{Hi,Wi,Ci} describe x, p.{Ho,Wo,Co} describe dy; p.M = R*S*Ci; p.Kg unused
  __shared__ __align__(16) float As[BK][BM + APAD];
  __shared__ __align__(16) float Bs[BK][BN + BPAD];
  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int64_t total_pix = (int64_t)p.N * p.Ho * p.Wo;
  const int64_t pix_lo = (int64_t)blockIdx.z * pix_per_split;
  int64_t pix_hi = pix_lo + pix_per_split;
  if (pix_hi > total_pix) pix_hi = total_pix;

  // A loader: 2 passes of 8 pixels x 32 lanes x 4 m-values
  const int am4 = (t & 31) * 4;
  const int akk = t >> 5;  // 0..7
  int a_r[4], a_s[4], a_c[4];
  bool a_ok[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int m = m0 + am4 + j;
    a_ok[j] = m < (int)p.M;
    int rs = a_ok[j] ? m / p.Ci : 0;
    a_c[j] = a_ok[j] ? m - rs * p.Ci : 0;
    a_r[j] = rs / p.S;
    a_s[j] = rs - a_r[j] * p.S;
  }
  // B loader: pixel t>>4, 4 filters at (t&15)*4
  const int bkk = t >> 4, bn4 = (t & 15) * 4;

  float areg[2][4], breg[4];
  auto load_tiles = [&](int64_t k0) {
    const int64_t hw = (int64_t)p.Ho * p.Wo;
#pragma unroll
    for (int ps = 0; ps < 2; ++ps) {
      int64_t pix = k0 + akk + ps * 8;
#pragma unroll
      for (int j = 0; j < 4; ++j) areg[ps][j] = 0.f;
      if (pix < pix_hi) {
        int n = (int)(pix / hw);
        int rem = (int)(pix - (int64_t)n * hw);
        int oh = rem / p.Wo, ow = rem - oh * p.Wo;
        if (VEC) {  // the 4 m-values share (r,s) and are channel-contiguous
          if (a_ok[0]) {
            int ih = oh * p.stride - p.pad + a_r[0], iw = ow * p.stride - p.pad + a_s[0];
            if (ih >= 0 && ih < p.Hi && iw >= 0 && iw < p.Wi)
              Vec4<T>::load(x + (((int64_t)n * p.Hi + ih) * p.Wi + iw) * p.Ci + a_c[0], areg[ps]);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (a_ok[j]) {
              int ih = oh * p.stride - p.pad + a_r[j], iw = ow * p.stride - p.pad + a_s[j];
              if (ih >= 0 && ih < p.Hi && iw >= 0 && iw < p.Wi)
                areg[ps][j] = to_f<T>(x[(((int64_t)n * p.Hi + ih) * p.Wi + iw) * p.Ci + a_c[j]]);
            }
          }
        }
      }
    }
    int64_t pix = k0 + bkk;
#pragma unroll
    for (int j = 0; j < 4; ++j) breg[j] = 0.f;
    if (pix < pix_hi) {
      if (n0 + bn4 + 3 < p.Co && (p.Co & 3) == 0) {
        Vec4<T>::load(dy + pix * p.Co + n0 + bn4, breg);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (n0 + bn4 + j < p.Co) breg[j] = to_f<T>(dy[pix * p.Co + n0 + bn4 + j]);
      }
    }
  };
  auto store_tiles = [&]() {
#pragma unroll
    for (int ps = 0; ps < 2; ++ps)
      *reinterpret_cast<float4*>(&As[akk + ps * 8][am4]) = make_float4(areg[ps][0], areg[ps][1], areg[ps][2], areg[ps][3]);
    *reinterpret_cast<float4*>(&Bs[bkk][bn4]) = make_float4(breg[0], breg[1], breg[2], breg[3]);
  };

  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  if (pix_lo < pix_hi) {
    load_tiles(pix_lo);
    store_tiles();
    __syncthreads();
    for (int64_t k0 = pix_lo; k0 < pix_hi; k0 += BK) {
      bool more = k0 + BK < pix_hi;
      if (more) load_tiles(k0 + BK);
      mma_tile(As, Bs, ty, tx, acc);
      __syncthreads();
      if (more) {
        store_tiles();
        __syncthreads();
      }
    }
  }
  // out[z][co][m], m contiguous: 8 consecutive m per thread per filter
  float* o = out + (int64_t)blockIdx.z * p.Co * p.M;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int co = n0 + tx * 4 + j;
    if (co >= p.Co) continue;
    int m = m0 + ty * 8;
    float* dst = o + (int64_t)co * p.M + m;
    if (m + 7 < (int)p.M && (p.M & 3) == 0) {
      *reinterpret_cast<float4*>(dst) = make_float4(acc[0][j], acc[1][j], acc[2][j], acc[3][j]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(acc[4][j], acc[5][j], acc[6][j], acc[7][j]);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (m + i < (int)p.M) dst[i] = acc[i][j];
    }
  }
}

__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ part, float* __restrict__ out,
                                                            int64_t n, int splits) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int z = 0; z < splits; ++z) s += part[(int64_t)z * n + i];  // fixed order: deterministic
    out[i] = s;
  }
}

// filter re-layout: KCRS fp32 (torch) -> KRSC or CRSK in T
template <typename T>
__global__ void filter_relayout_kernel(const float* __restrict__ w, T* __restrict__ o, int K, int C, int R, int S,
                                       int64_t sk, int64_t sc, int64_t sr, int64_t ss, int to_crsk) {
  int64_t n = (int64_t)K * C * R * S;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    // i indexes the OUTPUT layout
    int64_t rem = i;
    int k, c, r, s;
    if (!to_crsk) { c = rem % C; rem /= C; s = rem % S; rem /= S; r = rem % R; k = (int)(rem / R); }
    else { k = rem % K; rem /= K; s = rem % S; rem /= S; r = rem % R; c = (int)(rem / R); }
    o[i] = from_f<T>(w[k * sk + c * sc + r * sr + s * ss]);
  }
}

// every convolution's bf16 / fp32 KRSC and CRSK copies in ONE launch: blockIdx.y selects a table entry (the table lives
// in device memory at a fixed address, so the launch can be replayed from a CUDA graph after each optimizer update)
__global__ void __launch_bounds__(256) filter_relayout_batched_kernel(const iit_relayout_t* __restrict__ table) {
  const iit_relayout_t e = table[blockIdx.y];
  const uint32_t n = (uint32_t)e.k * e.c * e.r * e.s;            // 32-bit index arithmetic (host checks n < 2^31)
  const float* __restrict__ w = static_cast<const float*>(e.w);
  const uint32_t C = e.c, K = e.k, R = e.r, S = e.s;
  if (!e.to_crsk && e.dtype == IIT_BF16 && e.out_pitch == 0 && e.stride[1] == 1 && e.stride[3] == (int64_t)C &&
      e.stride[2] == (int64_t)S * C && e.stride[0] == (int64_t)R * S * C && (n & 3) == 0 &&
      ((reinterpret_cast<uintptr_t>(e.w) | reinterpret_cast<uintptr_t>(e.out)) & 15) == 0) {
    // channels-last master filter -> KRSC: the same order, a pure fp32 -> bf16 conversion (no index arithmetic)
    uint2* __restrict__ out = static_cast<uint2*>(e.out);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n / 4; i += gridDim.x * blockDim.x) {
      const float4 v = reinterpret_cast<const float4*>(w)[i];
      __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
      out[i] = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
    }
    return;
  }
  if (e.to_crsk && e.stride[1] == 1 && e.dtype == IIT_BF16 && e.out_pitch == 0) {
    // channels-last master filter -> CRSK: a [k][c] -> [c][k] transposition per tap. Through a 32 x 33 shared tile both the
    // fp32 reads (32 consecutive c) and the bf16 writes (32 consecutive k) are contiguous; element by element the reads were
    // R*S*C floats apart (157 us per refresh of all ResNet-18 filters).
    __shared__ float tile[32][33];
    const uint32_t tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const uint32_t tk = (K + 31) / 32, tc = (C + 31) / 32, ntiles = R * S * tk * tc;
    __nv_bfloat16* __restrict__ out = static_cast<__nv_bfloat16*>(e.out);
    for (uint32_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
      const uint32_t ct = t % tc, kt = (t / tc) % tk, rs = t / (tc * tk), r = rs / S, s2 = rs - r * S;
      const int64_t tap = (int64_t)r * e.stride[2] + (int64_t)s2 * e.stride[3];
#pragma unroll
      for (uint32_t j = 0; j < 4; ++j) {
        const uint32_t k = kt * 32 + ty + 8 * j, c = ct * 32 + tx;
        if (k < K && c < C) tile[ty + 8 * j][tx] = w[(int64_t)k * e.stride[0] + c + tap];
      }
      __syncthreads();
#pragma unroll
      for (uint32_t j = 0; j < 4; ++j) {
        const uint32_t c = ct * 32 + ty + 8 * j, k = kt * 32 + tx;
        if (k < K && c < C) out[((int64_t)(c * R + r) * S + s2) * K + k] = __float2bfloat16_rn(tile[tx][ty + 8 * j]);
      }
      __syncthreads();
    }
    return;
  }
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    uint32_t rem = i;   // i indexes the OUTPUT layout
    uint32_t k, c, r, s;
    if (!e.to_crsk) { c = rem % C; rem /= C; s = rem % S; rem /= S; r = rem % R; k = rem / R; }
    else { k = rem % K; rem /= K; s = rem % S; rem /= S; r = rem % R; c = rem / R; }
    const float v = w[(int64_t)k * e.stride[0] + (int64_t)c * e.stride[1] + (int64_t)r * e.stride[2] + (int64_t)s * e.stride[3]];
    int64_t o = i;
    if (e.out_pitch != 0) {   // rows of the output (outer index) are out_pitch elements apart
      const uint32_t inner = e.to_crsk ? R * S * K : R * S * C, outer = e.to_crsk ? c : k;
      o = (int64_t)outer * e.out_pitch + (i - outer * inner);
    }
    if (e.dtype == IIT_BF16) static_cast<__nv_bfloat16*>(e.out)[o] = __float2bfloat16_rn(v);
    else static_cast<float*>(e.out)[o] = v;
  }
}

__global__ void filter_grad_kernel(const float* __restrict__ g_krsc, float* __restrict__ g, int K, int C, int R, int S) {
  int64_t n = (int64_t)K * C * R * S;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t rem = i;  // i indexes KCRS
    int s = rem % S; rem /= S;
    int r = rem % R; rem /= R;
    int c = rem % C;
    int k = (int)(rem / C);
    g[i] = g_krsc[(((int64_t)k * R + r) * S + s) * C + c];
  }
}

static int check_conv(const iit_conv_t* d, const char* name) {
  IIT_REQUIRE(d, "%s: null descriptor", name);
  IIT_REQUIRE(d->n > 0 && d->h > 0 && d->w > 0 && d->c > 0 && d->k > 0 && d->r > 0 && d->s > 0 && d->stride > 0 && d->pad >= 0,
              "%s: non-positive dimension", name);
  int p = (d->h + 2 * d->pad - d->r) / d->stride + 1, q = (d->w + 2 * d->pad - d->s) / d->stride + 1;
  IIT_REQUIRE(p == d->p && q == d->q, "%s: output size mismatch: got %dx%d, expected %dx%d", name, d->p, d->q, p, q);
  return 0;
}

static int wgrad_splits(const iit_conv_t* d) {
  int64_t M = (int64_t)d->r * d->s * d->c;
  int64_t tiles = ((M + BM - 1) / BM) * ((d->k + BN - 1) / BN);
  int64_t pix = (int64_t)d->n * d->p * d->q;
  int64_t want = (2 * kNumSMs + tiles - 1) / tiles;
  int64_t maxs = (pix + BK * 8 - 1) / (BK * 8);
  if (want > maxs) want = maxs;
  if (want < 1) want = 1;
  if (want > 512) want = 512;
  return (int)want;
}

}  // namespace iit

using namespace iit;

extern "C" int iit_conv2d_fprop(const void* x, const void* w, void* y, const iit_conv_t* d, int dtype, double* bn_sum,
                                double* bn_sqsum, void* stream) {
  if (check_conv(d, "iit_conv2d_fprop")) return 1;
  IIT_REQUIRE(x && w && y, "iit_conv2d_fprop: null pointer");
  IIT_REQUIRE((bn_sum == nullptr) == (bn_sqsum == nullptr), "iit_conv2d_fprop: bn_sum and bn_sqsum go together");
  ConvP p{d->n, d->h, d->w, d->c, d->p, d->q, d->k, d->r, d->s, d->stride, d->pad, (int64_t)d->n * d->p * d->q, d->r * d->s * d->c};
  dim3 grid((unsigned)((p.M + BM - 1) / BM), (unsigned)((p.Co + BN - 1) / BN));
  cudaStream_t st = as_stream(stream);
  bool vec = (d->c % 16) == 0;
  IIT_DISPATCH_DTYPE(dtype, T, {
    if (vec) conv_gemm_kernel<T, false, true><<<grid, 256, 0, st>>>((const T*)x, (const T*)w, (T*)y, p, bn_sum, bn_sqsum);
    else conv_gemm_kernel<T, false, false><<<grid, 256, 0, st>>>((const T*)x, (const T*)w, (T*)y, p, bn_sum, bn_sqsum);
  })
  IIT_LAUNCH_CHECK("iit_conv2d_fprop");
  return 0;
}

extern "C" int iit_conv2d_dgrad(const void* dy, const void* w_crsk, void* dx, const iit_conv_t* d, int dtype, void* stream) {
  if (check_conv(d, "iit_conv2d_dgrad")) return 1;
  IIT_REQUIRE(dy && w_crsk && dx, "iit_conv2d_dgrad: null pointer");
  // gathered tensor = dy [n,p,q,k]; produced = dx [n,h,w,c]
  ConvP p{d->n, d->p, d->q, d->k, d->h, d->w, d->c, d->r, d->s, d->stride, d->pad, (int64_t)d->n * d->h * d->w, d->r * d->s * d->k};
  dim3 grid((unsigned)((p.M + BM - 1) / BM), (unsigned)((p.Co + BN - 1) / BN));
  cudaStream_t st = as_stream(stream);
  bool vec = (d->k % 16) == 0;
  IIT_DISPATCH_DTYPE(dtype, T, {
    if (vec) conv_gemm_kernel<T, true, true><<<grid, 256, 0, st>>>((const T*)dy, (const T*)w_crsk, (T*)dx, p, nullptr, nullptr);
    else conv_gemm_kernel<T, true, false><<<grid, 256, 0, st>>>((const T*)dy, (const T*)w_crsk, (T*)dx, p, nullptr, nullptr);
  })
  IIT_LAUNCH_CHECK("iit_conv2d_dgrad");
  return 0;
}

extern "C" size_t iit_conv2d_wgrad_scratch_bytes(const iit_conv_t* d) {
  if (!d) return 0;
  int splits = wgrad_splits(d);
  if (splits <= 1) return 0;
  return (size_t)splits * d->k * d->r * d->s * d->c * sizeof(float);
}

extern "C" int iit_conv2d_wgrad(const void* dy, const void* x, float* dw, const iit_conv_t* d, int dtype, void* scratch,
                                size_t scratch_bytes, void* stream) {
  if (check_conv(d, "iit_conv2d_wgrad")) return 1;
  IIT_REQUIRE(dy && x && dw, "iit_conv2d_wgrad: null pointer");
  int splits = wgrad_splits(d);
  size_t need = iit_conv2d_wgrad_scratch_bytes(d);
  IIT_REQUIRE(splits == 1 || (scratch && scratch_bytes >= need), "iit_conv2d_wgrad: scratch too small: need %zu bytes, got %zu",
              need, scratch_bytes);
  ConvP p{d->n, d->h, d->w, d->c, d->p, d->q, d->k, d->r, d->s, d->stride, d->pad, (int64_t)d->r * d->s * d->c, 0};
  int64_t pix = (int64_t)d->n * d->p * d->q;
  int64_t per = (pix + splits - 1) / splits;
  per = (per + BK - 1) / BK * BK;
  dim3 grid((unsigned)((p.M + BM - 1) / BM), (unsigned)((p.Co + BN - 1) / BN), (unsigned)splits);
  cudaStream_t st = as_stream(stream);
  float* target = splits == 1 ? dw : (float*)scratch;
  bool vec = (d->c % 4) == 0;
  IIT_DISPATCH_DTYPE(dtype, T, {
    if (vec) wgrad_kernel<T, true><<<grid, 256, 0, st>>>((const T*)x, (const T*)dy, target, p, per);
    else wgrad_kernel<T, false><<<grid, 256, 0, st>>>((const T*)x, (const T*)dy, target, p, per);
  })
  IIT_LAUNCH_CHECK("iit_conv2d_wgrad");
  if (splits > 1) {
    int64_t n = (int64_t)p.Co * p.M;
    splitk_reduce_kernel<<<grid_for(n, 256), 256, 0, st>>>((const float*)scratch, dw, n, splits);
    IIT_LAUNCH_CHECK("iit_conv2d_wgrad(reduce)");
  }
  return 0;
}

extern "C" int iit_filter_relayout(const float* w, const int64_t* strides_kcrs, void* w_out, int k, int c, int r, int s,
                                   int to_crsk, int dtype, void* stream) {
  IIT_REQUIRE(w && w_out && strides_kcrs, "iit_filter_relayout: null pointer");
  int64_t n = (int64_t)k * c * r * s;
  cudaStream_t st = as_stream(stream);
  IIT_DISPATCH_DTYPE(dtype, T, {
    filter_relayout_kernel<T><<<grid_for(n, 256), 256, 0, st>>>(w, (T*)w_out, k, c, r, s, strides_kcrs[0], strides_kcrs[1],
                                                               strides_kcrs[2], strides_kcrs[3], to_crsk);
  })
  IIT_LAUNCH_CHECK("iit_filter_relayout");
  return 0;
}

extern "C" int iit_filter_relayout_batched(const iit_relayout_t* table_dev, int entries, int64_t max_elements, void* stream) {
  IIT_REQUIRE(table_dev && entries >= 1 && max_elements >= 1 && max_elements < (1ll << 31), "iit_filter_relayout_batched: bad arguments");
  int bx = (int)((max_elements + 256 * 8 - 1) / (256 * 8));
  if (bx > 256) bx = 256;
  if (bx < 1) bx = 1;
  filter_relayout_batched_kernel<<<dim3((unsigned)bx, (unsigned)entries), 256, 0, as_stream(stream)>>>(table_dev);
  IIT_LAUNCH_CHECK("iit_filter_relayout_batched");
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Weight gradient for callers that cannot allocate device memory.
//
// Some hosts of this library cannot hand the kernels a scratch buffer: a plug-in called through an FFI
// that only forwards the tensors it was given; a service that reserved all device memory at start-up and
// must not call cudaMalloc afterwards (it synchronises the whole device and may fail once memory is
// fragmented); a latency-sensitive loop where that implicit synchronisation is not acceptable. For them
// the split-K partial sums of the weight gradient live in a fixed array inside the library image instead.
// The price: every GPU that loads the library carries kStaticScratchBytes of it until the process exits,
// and problems whose partial sums do not fit are refused.
// ------------------------------------------------------------------------------------------------
namespace iit {
constexpr size_t kStaticScratchBytes = 8u << 20;  // 8 MiB
static __device__ __align__(256) unsigned char g_wgrad_static_scratch[kStaticScratchBytes];
// device address of the array, per device; atomic because two host threads may make their first call at the same time (both
// then resolve the same address: the look-up is idempotent)
static std::atomic<void*> g_static_scratch_addr[kMaxDevices];
}  // namespace iit

extern "C" int iit_conv2d_wgrad_noalloc(const void* dy, const void* x, float* dw, const iit_conv_t* d, int dtype,
                                        void* stream) {
  if (check_conv(d, "iit_conv2d_wgrad_noalloc")) return 1;
  IIT_REQUIRE(dy && x && dw, "iit_conv2d_wgrad_noalloc: null pointer");
  const size_t need = iit_conv2d_wgrad_scratch_bytes(d);
  IIT_REQUIRE(need <= kStaticScratchBytes,
              "iit_conv2d_wgrad_noalloc: this problem needs %zu bytes of split-K scratch but the built-in buffer holds %zu; "
              "use iit_conv2d_wgrad with caller-provided scratch (or a smaller batch)", need, (size_t)kStaticScratchBytes);
  int dev = -1;
  IIT_CUDA(cudaGetDevice(&dev));
  IIT_REQUIRE(dev >= 0 && dev < kMaxDevices, "iit_conv2d_wgrad_noalloc: device ordinal %d out of range", dev);
  void* scratch = g_static_scratch_addr[dev].load(std::memory_order_acquire);
  if (scratch == nullptr) {
    // One-time set-up for the current device: in host code the array's name is only a handle, the
    // address the kernels must be given comes from the runtime. Nothing is allocated here.
    void* p = nullptr;
    IIT_CUDA(cudaGetSymbolAddress(&p, g_wgrad_static_scratch));
    IIT_REQUIRE(p != nullptr, "iit_conv2d_wgrad_noalloc: could not resolve the built-in scratch buffer");
    g_static_scratch_addr[dev].store(p, std::memory_order_release);
    scratch = p;
  }
  // The partial sums need no clearing: every cell of the region a launch uses is written by the
  // gradient kernel before the reduction reads it.
  return iit_conv2d_wgrad(dy, x, dw, d, dtype, scratch, kStaticScratchBytes, stream);
}
