// (b) Low-level HookedTransformer (IOI task: 6 layers, 4 heads, d_model 64, d_head 16, d_mlp 3072, 16 tokens):
// the non-GEMM pieces. Dense contractions (QKV, O, MLP in/out, unembed) run on the convolution GEMM kernels
// (a linear layer is a 1x1 convolution over tokens); here are embedding gather / scatter-add, LayerNormPre,
// bias (+GELU) (+residual) epilogues, column sums for bias gradients and the causal attention core.
// With d_model = 64 every one of these is HBM/latency bound: rows are short, so one warp owns one row
// (coalesced 256-byte row accesses), and attention keeps a whole (sequence, head) in shared memory.
#include <algorithm>

#include "common.cuh"

namespace iit {

// ---------------------------------------------------------------------------------------------- embedding
template <typename T>
__global__ void __launch_bounds__(256) embed_fwd_kernel(const int64_t* __restrict__ tok, const float* __restrict__ WE,
                                                        const float* __restrict__ Wpos, T* __restrict__ out, int64_t ntok,
                                                        int S, int D, int64_t V) {
  const int64_t total = ntok * (D / 4);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / (D / 4);
    const int d4 = (int)(i - t * (D / 4)) * 4;
    int64_t id = tok[t];
    if (id < 0) id += V;
    const int pos = (int)(t % S);
    float4 e = *reinterpret_cast<const float4*>(WE + id * D + d4);
    float4 p = *reinterpret_cast<const float4*>(Wpos + (int64_t)pos * D + d4);
    float v[4] = {e.x + p.x, e.y + p.y, e.z + p.z, e.w + p.w};
    Vec4<T>::store(out + t * D + d4, v);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) embed_bwd_kernel(const int64_t* __restrict__ tok, const T* __restrict__ dres,
                                                        float* __restrict__ dWE, float* __restrict__ dWpos, int64_t ntok,
                                                        int S, int D, int64_t V) {
  const int64_t total = ntok * D;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / D;
    const int d = (int)(i - t * D);
    int64_t id = tok[t];
    if (id < 0) id += V;
    const float g = to_f<T>(dres[i]);
    atomicAdd(dWE + id * D + d, g);
    atomicAdd(dWpos + (int64_t)(t % S) * D + d, g);
  }
}

// ---------------------------------------------------------------------------------------------- LayerNormPre
// y = (x - mean) / sqrt(mean((x-mean)^2) + eps); one warp per row, D a multiple of 32 (<= 1024)
template <typename T, int PER>
__global__ void __launch_bounds__(256) lnpre_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, float* __restrict__ rstd,
                                                        int64_t rows, int D, float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < rows; r += nw) {
    float v[PER];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < PER; ++j) { v[j] = to_f<T>(x[r * D + lane + 32 * j]); s += v[j]; }
    const float mean = warp_sum(s) / (float)D;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < PER; ++j) { v[j] -= mean; q += v[j] * v[j]; }
    const float scale = sqrtf(warp_sum(q) / (float)D + eps);
#pragma unroll
    for (int j = 0; j < PER; ++j) y[r * D + lane + 32 * j] = from_f<T>(v[j] / scale);
    if (lane == 0 && rstd) rstd[r] = 1.f / scale;
  }
}
// dx = rstd * (dy - mean(dy) - y * mean(dy * y))
template <typename T, int PER>
__global__ void __launch_bounds__(256) lnpre_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                        const float* __restrict__ rstd, T* __restrict__ dx, int64_t rows, int D,
                                                        int accumulate) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < rows; r += nw) {
    float g[PER], yy[PER];
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      g[j] = to_f<T>(dy[r * D + lane + 32 * j]);
      yy[j] = to_f<T>(y[r * D + lane + 32 * j]);
      a += g[j]; b += g[j] * yy[j];
    }
    a = warp_sum(a) / (float)D;
    b = warp_sum(b) / (float)D;
    const float rs = rstd[r];
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      float o = rs * (g[j] - a - yy[j] * b);
      if (accumulate) o += to_f<T>(dx[r * D + lane + 32 * j]);
      dx[r * D + lane + 32 * j] = from_f<T>(o);
    }
  }
}

// ---------------------------------------------------------------------------------------------- bias / GELU / residual
__device__ __forceinline__ float gelu_new_f(float x) {
  const float k = 0.7978845608028654f;  // sqrt(2/pi)
  return 0.5f * x * (1.f + tanhf(k * (x + 0.044715f * x * x * x)));
}
__device__ __forceinline__ float gelu_new_grad(float x) {
  const float k = 0.7978845608028654f;
  const float u = k * (x + 0.044715f * x * x * x);
  const float th = tanhf(u);
  return 0.5f * (1.f + th) + 0.5f * x * (1.f - th * th) * k * (1.f + 3.f * 0.044715f * x * x);
}

// pre = x + bias (+ resid);  out = act(pre).  pre_out optional (the MLP's hook_pre / what GELU's backward needs)
template <typename T>
__global__ void __launch_bounds__(256) bias_act_kernel(const T* __restrict__ x, const float* __restrict__ bias,
                                                       const T* __restrict__ resid, T* __restrict__ pre_out, T* __restrict__ out,
                                                       int64_t n4, int N, int act) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = i * 4;
    const int c = (int)(e % N);
    float v[4];
    Vec4<T>::load(x + e, v);
    if (bias) {
      float4 b = *reinterpret_cast<const float4*>(bias + c);
      v[0] += b.x; v[1] += b.y; v[2] += b.z; v[3] += b.w;
    }
    if (resid) {
      float r[4];
      Vec4<T>::load(resid + e, r);
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] += r[j];
    }
    if (pre_out) Vec4<T>::store(pre_out + e, v);
    if (act == 1) {
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = gelu_new_f(v[j]);
    }
    Vec4<T>::store(out + e, v);
  }
}
// bf16 fast path: 16 bytes per thread and access, 32-bit index arithmetic (the 64-bit `e % N` of the kernel above costs
// more than the rest of the loop), tanh.approx (2^-11 relative, below bf16's 2^-9 rounding)
__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void unpack_bf16x8(const uint4& t, float (&v)[8]) {
  const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) { v[2 * i] = __uint_as_float(w[i] << 16); v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
}
__device__ __forceinline__ uint4 pack_bf16x8(const float (&v)[8]) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    w[i] = *reinterpret_cast<uint32_t*>(&h);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}
__global__ void __launch_bounds__(256) bias_act_bf16x8_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ bias,
                                                              const __nv_bfloat16* __restrict__ resid,
                                                              __nv_bfloat16* __restrict__ pre_out, __nv_bfloat16* __restrict__ out,
                                                              uint32_t n8, uint32_t N, int act) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += gridDim.x * blockDim.x) {
    const int64_t e = (int64_t)i * 8;
    float v[8];
    unpack_bf16x8(*reinterpret_cast<const uint4*>(x + e), v);
    if (bias) {
      const uint32_t c = (i * 8u) % N;        // n8 * 8 < 2^32 (host-checked)
      const float4 b0 = *reinterpret_cast<const float4*>(bias + c), b1 = *reinterpret_cast<const float4*>(bias + c + 4);
      v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w; v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
    }
    if (resid) {
      float r[8];
      unpack_bf16x8(*reinterpret_cast<const uint4*>(resid + e), r);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += r[j];
    }
    if (pre_out) *reinterpret_cast<uint4*>(pre_out + e) = pack_bf16x8(v);
    if (act == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float u = 0.7978845608028654f * (v[j] + 0.044715f * v[j] * v[j] * v[j]);
        v[j] = 0.5f * v[j] * (1.f + tanh_fast(u));
      }
    }
    *reinterpret_cast<uint4*>(out + e) = pack_bf16x8(v);
  }
}
__global__ void __launch_bounds__(256) gelu_bwd_bf16x8_kernel(const __nv_bfloat16* __restrict__ dpost,
                                                              const __nv_bfloat16* __restrict__ pre, __nv_bfloat16* __restrict__ dpre,
                                                              uint32_t n8) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += gridDim.x * blockDim.x) {
    const int64_t e = (int64_t)i * 8;
    float g[8], p[8];
    unpack_bf16x8(*reinterpret_cast<const uint4*>(dpost + e), g);
    unpack_bf16x8(*reinterpret_cast<const uint4*>(pre + e), p);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float k = 0.7978845608028654f, x = p[j];
      const float th = tanh_fast(k * (x + 0.044715f * x * x * x));
      g[j] *= 0.5f * (1.f + th) + 0.5f * x * (1.f - th * th) * k * (1.f + 3.f * 0.044715f * x * x);
    }
    *reinterpret_cast<uint4*>(dpre + e) = pack_bf16x8(g);
  }
}
static bool aligned16(const void* a, const void* b = nullptr, const void* c = nullptr, const void* d = nullptr) {
  return ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c) |
           reinterpret_cast<uintptr_t>(d)) & 15) == 0;
}
// same, element by element, for widths that are not a multiple of 4 (the 50257-wide unembed bias)
template <typename T>
__global__ void __launch_bounds__(256) bias_act_scalar_kernel(const T* __restrict__ x, const float* __restrict__ bias,
                                                              const T* __restrict__ resid, T* __restrict__ pre_out,
                                                              T* __restrict__ out, int64_t n, int N, int act) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = to_f<T>(x[i]);
    if (bias) v += bias[i % N];
    if (resid) v += to_f<T>(resid[i]);
    if (pre_out) pre_out[i] = from_f<T>(v);
    if (act == 1) v = gelu_new_f(v);
    out[i] = from_f<T>(v);
  }
}
template <typename T>
__global__ void __launch_bounds__(256) gelu_bwd_kernel(const T* __restrict__ dpost, const T* __restrict__ pre, T* __restrict__ dpre,
                                                       int64_t n4) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float g[4], p[4];
    Vec4<T>::load(dpost + i * 4, g);
    Vec4<T>::load(pre + i * 4, p);
#pragma unroll
    for (int j = 0; j < 4; ++j) g[j] *= gelu_new_grad(p[j]);
    Vec4<T>::store(dpre + i * 4, g);
  }
}

// out[c] += sum_r x[r, c]  (bias gradients). Block = 32 columns x 8 row lanes.
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x, float* __restrict__ out, int64_t rows, int64_t N) {
  __shared__ float s[8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int64_t c = (int64_t)blockIdx.x * 32 + cx;
  float acc = 0.f;
  if (c < N)
    for (int64_t r = (int64_t)blockIdx.y * 8 + ry; r < rows; r += (int64_t)gridDim.y * 8) acc += to_f<T>(x[r * N + c]);
  s[ry][cx] = acc;
  __syncthreads();
  if (ry == 0 && c < N) {
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += s[j][cx];
    atomicAdd(out + c, t);
  }
}

// bf16, N % 8 == 0: a thread owns 8 consecutive columns (16-byte loads, four rows in flight); block = 16 column packs x 16 row
// lanes, shared-memory reduction over the row lanes, one atomic per column and block
__global__ void __launch_bounds__(256) colsum_bf16x8_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out,
                                                            int64_t rows, int N) {
  __shared__ float s[16][16 * 8 + 1];
  const int cp = threadIdx.x & 15, ry = threadIdx.x >> 4;
  const int c = (blockIdx.x * 16 + cp) * 8;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c < N) {
    const int64_t step = (int64_t)gridDim.y * 16;
    int64_t r = (int64_t)blockIdx.y * 16 + ry;
    for (; r + 3 * step < rows; r += 4 * step) {
      uint4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) t[u] = *reinterpret_cast<const uint4*>(x + (r + u * step) * N + c);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float v[8];
        unpack_bf16x8(t[u], v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += v[j];
      }
    }
    for (; r < rows; r += step) {
      float v[8];
      unpack_bf16x8(*reinterpret_cast<const uint4*>(x + r * N + c), v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) s[ry][cp * 8 + j] = acc[j];
  __syncthreads();
  if (threadIdx.x < 128) {
    const int col = blockIdx.x * 128 + threadIdx.x;
    if (col < N) {
      float t = 0.f;
#pragma unroll
      for (int j = 0; j < 16; ++j) t += s[j][threadIdx.x];
      atomicAdd(out + col, t);
    }
  }
}

// ---------------------------------------------------------------------------------------------- attention core
// qkv: [B, S, 3, H, DH] (one fused projection), z: [B, S, H, DH]. Causal, scores / sqrt(DH).
// One warp per (batch, head); lane i < S owns query row i. S <= 32, DH <= 32.
template <typename T, int S_MAX, int DH_MAX>
__global__ void __launch_bounds__(128) attn_fwd_kernel(const T* __restrict__ qkv, T* __restrict__ z, int64_t BH, int S, int H,
                                                       int DH, float scale) {
  // a (batch, head) unit needs S lanes: with S_MAX = 16 a warp serves two units (lanes 0-15 and 16-31)
  constexpr int UPW = 32 / S_MAX, UPB = 4 * UPW;
  __shared__ float sk[UPB][S_MAX][DH_MAX + 1], sv[UPB][S_MAX][DH_MAX + 1];
  const int lane = threadIdx.x & 31, row = lane % S_MAX;
  const int w = (threadIdx.x >> 5) * UPW + lane / S_MAX;       // unit slot within the block
  for (int64_t bh0 = (int64_t)blockIdx.x * UPB; bh0 < BH; bh0 += (int64_t)gridDim.x * UPB) {
    const int64_t bh = bh0 + w;
    const bool live = bh < BH;
    const int64_t b = live ? bh / H : 0;
    const int h = live ? (int)(bh - b * H) : 0;
    const T* base = qkv + b * (int64_t)S * 3 * H * DH;
    if (live)
      for (int i = row; i < S * DH; i += S_MAX) {
        const int s = i / DH, d = i - s * DH;
        sk[w][s][d] = to_f<T>(base[((int64_t)s * 3 + 1) * H * DH + h * DH + d]);
        sv[w][s][d] = to_f<T>(base[((int64_t)s * 3 + 2) * H * DH + h * DH + d]);
      }
    __syncwarp();
    if (live && row < S) {
      float q[DH_MAX], o[DH_MAX];
#pragma unroll
      for (int d = 0; d < DH_MAX; ++d) {
        q[d] = d < DH ? to_f<T>(base[((int64_t)row * 3) * H * DH + h * DH + d]) : 0.f;
        o[d] = 0.f;
      }
      float m = -INFINITY, l = 0.f;
      for (int j = 0; j <= row; ++j) {
        float sc = 0.f;
#pragma unroll
        for (int d = 0; d < DH_MAX; ++d) if (d < DH) sc = fmaf(q[d], sk[w][j][d], sc);
        sc *= scale;
        const float mn = fmaxf(m, sc);
        const float corr = __expf(m - mn), p = __expf(sc - mn);
        l = l * corr + p;
#pragma unroll
        for (int d = 0; d < DH_MAX; ++d) if (d < DH) o[d] = o[d] * corr + p * sv[w][j][d];
        m = mn;
      }
      T* zo = z + ((b * S + row) * H + h) * DH;
#pragma unroll
      for (int d = 0; d < DH_MAX; ++d) if (d < DH) zo[d] = from_f<T>(o[d] / l);
    }
    __syncwarp();
  }
}

// dqkv [B,S,3,H,DH] from dz, recomputing the pattern.
template <typename T, int S_MAX, int DH_MAX>
__global__ void __launch_bounds__(64) attn_bwd_kernel(const T* __restrict__ qkv, const T* __restrict__ dz, T* __restrict__ dqkv,
                                                       int64_t BH, int S, int H, int DH, float scale) {
  constexpr int UPW = 32 / S_MAX, UPB = 2 * UPW;     // units per warp as in the forward; two warps per block (shared memory)
  __shared__ float sq[UPB][S_MAX][DH_MAX + 1], sk[UPB][S_MAX][DH_MAX + 1], sv[UPB][S_MAX][DH_MAX + 1], sdz[UPB][S_MAX][DH_MAX + 1];
  __shared__ float sp[UPB][S_MAX][S_MAX + 1], sds[UPB][S_MAX][S_MAX + 1];
  const int lane = threadIdx.x & 31, row = lane % S_MAX;
  const int w = (threadIdx.x >> 5) * UPW + lane / S_MAX;
  for (int64_t bh0 = (int64_t)blockIdx.x * UPB; bh0 < BH; bh0 += (int64_t)gridDim.x * UPB) {
    const int64_t bh = bh0 + w;
    const bool live = bh < BH;
    const int64_t b = live ? bh / H : 0;
    const int h = live ? (int)(bh - b * H) : 0;
    const T* base = qkv + b * (int64_t)S * 3 * H * DH;
    T* dbase = dqkv + b * (int64_t)S * 3 * H * DH;
    if (live)
      for (int i = row; i < S * DH; i += S_MAX) {
        const int s = i / DH, d = i - s * DH;
        sq[w][s][d] = to_f<T>(base[((int64_t)s * 3 + 0) * H * DH + h * DH + d]);
        sk[w][s][d] = to_f<T>(base[((int64_t)s * 3 + 1) * H * DH + h * DH + d]);
        sv[w][s][d] = to_f<T>(base[((int64_t)s * 3 + 2) * H * DH + h * DH + d]);
        sdz[w][s][d] = to_f<T>(dz[((b * S + s) * H + h) * DH + d]);
      }
    __syncwarp();
    if (live && row < S) {  // phase 1: row i of the pattern P and of dS
      const int i = row;
      // scores -> pattern row (kept in shared memory: sp), dP row (kept in sds until rescaled)
      float m = -INFINITY;
      for (int j = 0; j <= i; ++j) {
        float a = 0.f;
        for (int d = 0; d < DH; ++d) a = fmaf(sq[w][i][d], sk[w][j][d], a);
        a *= scale;
        sp[w][i][j] = a;
        m = fmaxf(m, a);
      }
      float l = 0.f;
      for (int j = 0; j <= i; ++j) { const float e = __expf(sp[w][i][j] - m); sp[w][i][j] = e; l += e; }
      float dot = 0.f;
      for (int j = 0; j <= i; ++j) {
        const float p = sp[w][i][j] / l;
        float a = 0.f;
        for (int d = 0; d < DH; ++d) a = fmaf(sdz[w][i][d], sv[w][j][d], a);
        sds[w][i][j] = a;
        dot += p * a;
        sp[w][i][j] = p;
      }
      for (int j = 0; j < S; ++j) {
        if (j <= i) sds[w][i][j] = sp[w][i][j] * (sds[w][i][j] - dot) * scale;
        else { sds[w][i][j] = 0.f; sp[w][i][j] = 0.f; }
      }
      // dq_i = sum_j dS_ij k_j
      for (int d = 0; d < DH; ++d) {
        float a = 0.f;
        for (int j = 0; j <= i; ++j) a = fmaf(sds[w][i][j], sk[w][j][d], a);
        dbase[((int64_t)i * 3 + 0) * H * DH + h * DH + d] = from_f<T>(a);
      }
    }
    __syncwarp();
    if (live && row < S) {  // phase 2: column j: dk_j = sum_i dS_ij q_i, dv_j = sum_i P_ij dz_i
      const int j = row;
      for (int d = 0; d < DH; ++d) {
        float a = 0.f, c = 0.f;
        for (int i = j; i < S; ++i) { a = fmaf(sds[w][i][j], sq[w][i][d], a); c = fmaf(sp[w][i][j], sdz[w][i][d], c); }
        dbase[((int64_t)j * 3 + 1) * H * DH + h * DH + d] = from_f<T>(a);
        dbase[((int64_t)j * 3 + 2) * H * DH + h * DH + d] = from_f<T>(c);
      }
    }
    __syncwarp();
  }
}

// ---- the IOI model's attention (16 tokens, d_head 16, bf16): the same arithmetic with 16-byte loads and stores ------------
// A token's q / k / v row of one head is 32 contiguous bytes of qkv [B, S, 3, H, 16]; lane `row` of a half-warp owns token
// `row` of its (batch, head) unit: it loads its own rows with two 16-byte loads each (the general kernels issue sixteen 2-byte
// loads per row) and writes its 16 outputs with two 16-byte stores.
__device__ __forceinline__ void ld_row16(const __nv_bfloat16* p, float (&v)[16]) {
  float a[8], b[8];
  unpack_bf16x8(*reinterpret_cast<const uint4*>(p), a);
  unpack_bf16x8(*reinterpret_cast<const uint4*>(p + 8), b);
#pragma unroll
  for (int i = 0; i < 8; ++i) { v[i] = a[i]; v[8 + i] = b[i]; }
}
__device__ __forceinline__ void st_row16(__nv_bfloat16* p, const float (&v)[16]) {
  float a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = v[i]; b[i] = v[8 + i]; }
  *reinterpret_cast<uint4*>(p) = pack_bf16x8(a);
  *reinterpret_cast<uint4*>(p + 8) = pack_bf16x8(b);
}

__global__ void __launch_bounds__(128) attn_fwd_s16_bf16_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ z,
                                                                int64_t BH, int H, float scale) {
  constexpr int S = 16, DH = 16, UPB = 8, PITCH = 20;      // two units per warp, four warps; rows padded to 80 bytes
  __shared__ __align__(16) float sk[UPB][S][PITCH], sv[UPB][S][PITCH];
  const int lane = threadIdx.x & 31, row = lane & 15;
  const int w = (threadIdx.x >> 5) * 2 + (lane >> 4);
  for (int64_t bh0 = (int64_t)blockIdx.x * UPB; bh0 < BH; bh0 += (int64_t)gridDim.x * UPB) {
    const int64_t bh = bh0 + w;
    const bool live = bh < BH;
    const int64_t b = live ? bh / H : 0;
    const int h = live ? (int)(bh - b * H) : 0;
    const __nv_bfloat16* tok = qkv + ((b * S + row) * 3) * (int64_t)H * DH + h * DH;     // this lane's token: q row
    float q[DH], o[DH];
    if (live) {
      float kv[DH];
      ld_row16(tok, q);
      ld_row16(tok + (int64_t)H * DH, kv);
#pragma unroll
      for (int d = 0; d < DH; d += 4) *reinterpret_cast<float4*>(&sk[w][row][d]) = make_float4(kv[d], kv[d + 1], kv[d + 2], kv[d + 3]);
      ld_row16(tok + 2 * (int64_t)H * DH, kv);
#pragma unroll
      for (int d = 0; d < DH; d += 4) *reinterpret_cast<float4*>(&sv[w][row][d]) = make_float4(kv[d], kv[d + 1], kv[d + 2], kv[d + 3]);
    }
    __syncwarp();
    if (live) {
#pragma unroll
      for (int d = 0; d < DH; ++d) o[d] = 0.f;
      float m = -INFINITY, l = 0.f;
      for (int j = 0; j <= row; ++j) {
        float sc = 0.f;
#pragma unroll
        for (int d = 0; d < DH; ++d) sc = fmaf(q[d], sk[w][j][d], sc);
        sc *= scale;
        const float mn = fmaxf(m, sc);
        const float corr = __expf(m - mn), pr = __expf(sc - mn);
        l = l * corr + pr;
#pragma unroll
        for (int d = 0; d < DH; ++d) o[d] = o[d] * corr + pr * sv[w][j][d];
        m = mn;
      }
      const float inv = 1.f / l;
#pragma unroll
      for (int d = 0; d < DH; ++d) o[d] *= inv;
      st_row16(z + ((b * S + row) * H + h) * DH, o);
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(64) attn_bwd_s16_bf16_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ dz,
                                                               __nv_bfloat16* __restrict__ dqkv, int64_t BH, int H, float scale) {
  constexpr int S = 16, DH = 16, UPB = 4, PITCH = 20;
  __shared__ __align__(16) float sq[UPB][S][PITCH], sk[UPB][S][PITCH], sv[UPB][S][PITCH], sdz[UPB][S][PITCH];
  __shared__ float sp[UPB][S][S + 1], sds[UPB][S][S + 1];
  const int lane = threadIdx.x & 31, row = lane & 15;
  const int w = (threadIdx.x >> 5) * 2 + (lane >> 4);
  for (int64_t bh0 = (int64_t)blockIdx.x * UPB; bh0 < BH; bh0 += (int64_t)gridDim.x * UPB) {
    const int64_t bh = bh0 + w;
    const bool live = bh < BH;
    const int64_t b = live ? bh / H : 0;
    const int h = live ? (int)(bh - b * H) : 0;
    const int64_t tok_off = ((b * S + row) * 3) * (int64_t)H * DH + h * DH;
    float q[DH], g[DH];            // this lane's own query row and dz row
    if (live) {
      float t[DH];
      ld_row16(qkv + tok_off, q);
#pragma unroll
      for (int d = 0; d < DH; d += 4) *reinterpret_cast<float4*>(&sq[w][row][d]) = make_float4(q[d], q[d + 1], q[d + 2], q[d + 3]);
      ld_row16(qkv + tok_off + (int64_t)H * DH, t);
#pragma unroll
      for (int d = 0; d < DH; d += 4) *reinterpret_cast<float4*>(&sk[w][row][d]) = make_float4(t[d], t[d + 1], t[d + 2], t[d + 3]);
      ld_row16(qkv + tok_off + 2 * (int64_t)H * DH, t);
#pragma unroll
      for (int d = 0; d < DH; d += 4) *reinterpret_cast<float4*>(&sv[w][row][d]) = make_float4(t[d], t[d + 1], t[d + 2], t[d + 3]);
      ld_row16(dz + ((b * S + row) * H + h) * DH, g);
#pragma unroll
      for (int d = 0; d < DH; d += 4) *reinterpret_cast<float4*>(&sdz[w][row][d]) = make_float4(g[d], g[d + 1], g[d + 2], g[d + 3]);
    }
    __syncwarp();
    if (live) {   // phase 1: row i of the pattern P and of dS; dq_i
      const int i = row;
      float m = -INFINITY;
      for (int j = 0; j <= i; ++j) {
        float a = 0.f;
#pragma unroll
        for (int d = 0; d < DH; ++d) a = fmaf(q[d], sk[w][j][d], a);
        a *= scale;
        sp[w][i][j] = a;
        m = fmaxf(m, a);
      }
      float l = 0.f;
      for (int j = 0; j <= i; ++j) { const float e = __expf(sp[w][i][j] - m); sp[w][i][j] = e; l += e; }
      const float inv = 1.f / l;
      float dot = 0.f;
      for (int j = 0; j <= i; ++j) {
        const float pr = sp[w][i][j] * inv;
        float a = 0.f;
#pragma unroll
        for (int d = 0; d < DH; ++d) a = fmaf(g[d], sv[w][j][d], a);
        sds[w][i][j] = a;
        dot += pr * a;
        sp[w][i][j] = pr;
      }
      float dq[DH];
#pragma unroll
      for (int d = 0; d < DH; ++d) dq[d] = 0.f;
      for (int j = 0; j < S; ++j) {
        if (j <= i) {
          const float ds = sp[w][i][j] * (sds[w][i][j] - dot) * scale;
          sds[w][i][j] = ds;
#pragma unroll
          for (int d = 0; d < DH; ++d) dq[d] = fmaf(ds, sk[w][j][d], dq[d]);
        } else { sds[w][i][j] = 0.f; sp[w][i][j] = 0.f; }
      }
      st_row16(dqkv + tok_off, dq);
    }
    __syncwarp();
    if (live) {   // phase 2: column j: dk_j = sum_i dS_ij q_i, dv_j = sum_i P_ij dz_i
      const int j = row;
      float dk[DH], dv[DH];
#pragma unroll
      for (int d = 0; d < DH; ++d) { dk[d] = 0.f; dv[d] = 0.f; }
      for (int i = j; i < S; ++i) {
        const float ds = sds[w][i][j], pr = sp[w][i][j];
#pragma unroll
        for (int d = 0; d < DH; ++d) { dk[d] = fmaf(ds, sq[w][i][d], dk[d]); dv[d] = fmaf(pr, sdz[w][i][d], dv[d]); }
      }
      st_row16(dqkv + tok_off + (int64_t)H * DH, dk);
      st_row16(dqkv + tok_off + 2 * (int64_t)H * DH, dv);
    }
    __syncwarp();
  }
}

// ---- LayerNormPre for d_model = 64 in bf16: eight lanes per row (16-byte accesses), four rows per warp ----------------------
__device__ __forceinline__ float sum8(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  return v;
}
__global__ void __launch_bounds__(256) lnpre_fwd_d64_bf16_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                                                 float* __restrict__ rstd, int64_t rows, float eps) {
  const int64_t r = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 3;      // grid covers ceil(rows / 32) blocks exactly
  const int part = threadIdx.x & 7;
  const bool live = r < rows;
  float v[8];
  if (live) unpack_bf16x8(*reinterpret_cast<const uint4*>(x + r * 64 + part * 8), v);
  else {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += v[j];
  const float mean = sum8(s) * (1.f / 64.f);
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) { v[j] -= mean; q += v[j] * v[j]; }
  const float scale = sqrtf(sum8(q) * (1.f / 64.f) + eps);
  if (!live) return;
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = v[j] / scale;
  *reinterpret_cast<uint4*>(y + r * 64 + part * 8) = pack_bf16x8(v);
  if (part == 0 && rstd) rstd[r] = 1.f / scale;
}
__global__ void __launch_bounds__(256) lnpre_bwd_d64_bf16_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                                                                 const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dx,
                                                                 int64_t rows, int accumulate) {
  const int64_t r = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 3;
  const int part = threadIdx.x & 7;
  const bool live = r < rows;
  float g[8], yy[8], old[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) old[j] = 0.f;
  if (live) {
    unpack_bf16x8(*reinterpret_cast<const uint4*>(dy + r * 64 + part * 8), g);
    unpack_bf16x8(*reinterpret_cast<const uint4*>(y + r * 64 + part * 8), yy);
    if (accumulate) unpack_bf16x8(*reinterpret_cast<const uint4*>(dx + r * 64 + part * 8), old);
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) { g[j] = 0.f; yy[j] = 0.f; }
  }
  float a = 0.f, b = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) { a += g[j]; b += g[j] * yy[j]; }
  a = sum8(a) * (1.f / 64.f);
  b = sum8(b) * (1.f / 64.f);
  if (!live) return;
  const float rs = rstd[r];
#pragma unroll
  for (int j = 0; j < 8; ++j) g[j] = rs * (g[j] - a - yy[j] * b) + old[j];
  *reinterpret_cast<uint4*>(dx + r * 64 + part * 8) = pack_bf16x8(g);
}

}  // namespace iit

using namespace iit;

extern "C" int iit_embed_fwd(const int64_t* tokens, const float* W_E, const float* W_pos, void* out, int dtype, int64_t ntok,
                             int seq, int d_model, int64_t d_vocab, void* stream) {
  IIT_REQUIRE(tokens && W_E && W_pos && out && d_model % 4 == 0 && seq >= 1, "iit_embed_fwd: bad arguments");
  IIT_DISPATCH_DTYPE(dtype, T, {
    embed_fwd_kernel<T><<<grid_for(ntok * (d_model / 4), 256), 256, 0, as_stream(stream)>>>(tokens, W_E, W_pos, (T*)out, ntok, seq,
                                                                                           d_model, d_vocab);
  })
  IIT_LAUNCH_CHECK("iit_embed_fwd");
  return 0;
}
extern "C" int iit_embed_bwd(const int64_t* tokens, const void* dresid, int dtype, float* dW_E, float* dW_pos, int64_t ntok,
                             int seq, int d_model, int64_t d_vocab, void* stream) {
  IIT_REQUIRE(tokens && dresid && dW_E && dW_pos, "iit_embed_bwd: null pointer");
  IIT_DISPATCH_DTYPE(dtype, T, {
    embed_bwd_kernel<T><<<grid_for(ntok * d_model, 256), 256, 0, as_stream(stream)>>>(tokens, (const T*)dresid, dW_E, dW_pos, ntok,
                                                                                     seq, d_model, d_vocab);
  })
  IIT_LAUNCH_CHECK("iit_embed_bwd");
  return 0;
}

#define IIT_LN_DISPATCH(KERNEL, ...)                                                                   \
  switch (d / 32) {                                                                                    \
    case 1: KERNEL<T, 1><<<grid_for(rows, 8), 256, 0, st>>>(__VA_ARGS__); break;                       \
    case 2: KERNEL<T, 2><<<grid_for(rows, 8), 256, 0, st>>>(__VA_ARGS__); break;                       \
    case 4: KERNEL<T, 4><<<grid_for(rows, 8), 256, 0, st>>>(__VA_ARGS__); break;                       \
    case 8: KERNEL<T, 8><<<grid_for(rows, 8), 256, 0, st>>>(__VA_ARGS__); break;                       \
    case 16: KERNEL<T, 16><<<grid_for(rows, 8), 256, 0, st>>>(__VA_ARGS__); break;                     \
    case 24: KERNEL<T, 24><<<grid_for(rows, 8), 256, 0, st>>>(__VA_ARGS__); break;                     \
    case 32: KERNEL<T, 32><<<grid_for(rows, 8), 256, 0, st>>>(__VA_ARGS__); break;                     \
    default: iit::set_error("LayerNormPre: unsupported width %d", d); return 1;                        \
  }

extern "C" int iit_lnpre_fwd(const void* x, void* y, float* rstd, int dtype, int64_t rows, int d, float eps, void* stream) {
  IIT_REQUIRE(x && y && d % 32 == 0, "iit_lnpre_fwd: bad arguments (width must be a multiple of 32)");
  cudaStream_t st = as_stream(stream);
  if (dtype == IIT_BF16 && d == 64 && aligned16(x, y) && rows > 0) {
    lnpre_fwd_d64_bf16_kernel<<<(unsigned)((rows + 31) / 32), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, rstd, rows, eps);
    IIT_LAUNCH_CHECK("iit_lnpre_fwd");
    return 0;
  }
  IIT_DISPATCH_DTYPE(dtype, T, { IIT_LN_DISPATCH(lnpre_fwd_kernel, (const T*)x, (T*)y, rstd, rows, d, eps) })
  IIT_LAUNCH_CHECK("iit_lnpre_fwd");
  return 0;
}
static int lnpre_bwd_impl(const void* dy, const void* y, const float* rstd, void* dx, int dtype, int64_t rows, int d, void* stream,
                          int accumulate, const char* name) {
  IIT_REQUIRE(dy && y && rstd && dx && d % 32 == 0, "%s: bad arguments", name);
  cudaStream_t st = as_stream(stream);
  if (dtype == IIT_BF16 && d == 64 && aligned16(dy, y, dx) && rows > 0) {
    lnpre_bwd_d64_bf16_kernel<<<(unsigned)((rows + 31) / 32), 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, rstd,
                                                                          (__nv_bfloat16*)dx, rows, accumulate);
    IIT_LAUNCH_CHECK(name);
    return 0;
  }
  IIT_DISPATCH_DTYPE(dtype, T, { IIT_LN_DISPATCH(lnpre_bwd_kernel, (const T*)dy, (const T*)y, rstd, (T*)dx, rows, d, accumulate) })
  IIT_LAUNCH_CHECK(name);
  return 0;
}
extern "C" int iit_lnpre_bwd(const void* dy, const void* y, const float* rstd, void* dx, int dtype, int64_t rows, int d,
                             void* stream) {
  return lnpre_bwd_impl(dy, y, rstd, dx, dtype, rows, d, stream, 0, "iit_lnpre_bwd");
}
extern "C" int iit_lnpre_bwd_accumulate(const void* dy, const void* y, const float* rstd, void* dx, int dtype, int64_t rows, int d,
                                        void* stream) {
  return lnpre_bwd_impl(dy, y, rstd, dx, dtype, rows, d, stream, 1, "iit_lnpre_bwd_accumulate");
}

extern "C" int iit_bias_act_fwd(const void* x, const float* bias, const void* resid, void* pre_out, void* out, int dtype,
                                int64_t rows, int n, int act, void* stream) {
  IIT_REQUIRE(x && out && (act == 0 || act == 1), "iit_bias_act_fwd: bad arguments");
  if (n % 4 != 0) {
    IIT_DISPATCH_DTYPE(dtype, T, {
      bias_act_scalar_kernel<T><<<grid_for(rows * n, 1024), 256, 0, as_stream(stream)>>>((const T*)x, bias, (const T*)resid,
                                                                                        (T*)pre_out, (T*)out, rows * n, n, act);
    })
    IIT_LAUNCH_CHECK("iit_bias_act_fwd");
    return 0;
  }
  if (dtype == IIT_BF16 && n % 8 == 0 && rows * n < (1ll << 32) && aligned16(x, resid, pre_out, out) && aligned16(bias)) {
    const int64_t n8 = rows * n / 8;
    bias_act_bf16x8_kernel<<<grid_for(n8, 512), 256, 0, as_stream(stream)>>>(
        (const __nv_bfloat16*)x, bias, (const __nv_bfloat16*)resid, (__nv_bfloat16*)pre_out, (__nv_bfloat16*)out, (uint32_t)n8,
        (uint32_t)n, act);
    IIT_LAUNCH_CHECK("iit_bias_act_fwd");
    return 0;
  }
  int64_t n4 = rows * n / 4;
  IIT_DISPATCH_DTYPE(dtype, T, {
    bias_act_kernel<T><<<grid_for(n4, 512), 256, 0, as_stream(stream)>>>((const T*)x, bias, (const T*)resid, (T*)pre_out, (T*)out,
                                                                        n4, n, act);
  })
  IIT_LAUNCH_CHECK("iit_bias_act_fwd");
  return 0;
}
extern "C" int iit_gelu_bwd(const void* dpost, const void* pre, void* dpre, int dtype, int64_t n, void* stream) {
  IIT_REQUIRE(dpost && pre && dpre && n % 4 == 0, "iit_gelu_bwd: bad arguments");
  if (dtype == IIT_BF16 && n % 8 == 0 && n < (1ll << 32) && aligned16(dpost, pre, dpre)) {
    gelu_bwd_bf16x8_kernel<<<grid_for(n / 8, 512), 256, 0, as_stream(stream)>>>((const __nv_bfloat16*)dpost, (const __nv_bfloat16*)pre,
                                                                              (__nv_bfloat16*)dpre, (uint32_t)(n / 8));
    IIT_LAUNCH_CHECK("iit_gelu_bwd");
    return 0;
  }
  IIT_DISPATCH_DTYPE(dtype, T, {
    gelu_bwd_kernel<T><<<grid_for(n / 4, 512), 256, 0, as_stream(stream)>>>((const T*)dpost, (const T*)pre, (T*)dpre, n / 4);
  })
  IIT_LAUNCH_CHECK("iit_gelu_bwd");
  return 0;
}
// dpre = dpost * gelu_new'(h + bias) and, in the same pass, colsum[c] += sum_r dpre[r, c] (the input bias's gradient): the
// geometry of colsum_bf16x8_kernel (a thread owns 8 columns; 16 column packs x 16 row lanes per block; two rows in flight).
// `h` is the MLP input map's raw GEMM result: the pre-activation h + bias is never stored.
__global__ void __launch_bounds__(256) gelu_bwd_bias_colsum_kernel(const __nv_bfloat16* __restrict__ dpost,
                                                                   const __nv_bfloat16* __restrict__ h, const float* __restrict__ bias,
                                                                   __nv_bfloat16* __restrict__ dpre, float* __restrict__ colsum,
                                                                   int64_t rows, int N) {
  __shared__ float s[16][16 * 8 + 1];
  const int cp = threadIdx.x & 15, ry = threadIdx.x >> 4;
  const int c = (blockIdx.x * 16 + cp) * 8;
  float acc[8], b[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { acc[j] = 0.f; b[j] = 0.f; }
  if (c < N) {
    if (bias != nullptr) {
      const float4 b0 = *reinterpret_cast<const float4*>(bias + c), b1 = *reinterpret_cast<const float4*>(bias + c + 4);
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
    }
    const int64_t step = (int64_t)gridDim.y * 16;
    auto one = [&](const uint4& tg, const uint4& th, int64_t r) {
      float g[8], x[8];
      unpack_bf16x8(tg, g);
      unpack_bf16x8(th, x);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float k = 0.7978845608028654f, v = x[j] + b[j];
        const float t = tanh_fast(k * (v + 0.044715f * v * v * v));
        g[j] *= 0.5f * (1.f + t) + 0.5f * v * (1.f - t * t) * k * (1.f + 3.f * 0.044715f * v * v);
      }
      const uint4 o = pack_bf16x8(g);
      *reinterpret_cast<uint4*>(dpre + r * N + c) = o;
      unpack_bf16x8(o, g);          // the sums are of the values as stored
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += g[j];
    };
    int64_t r = (int64_t)blockIdx.y * 16 + ry;
    for (; r + step < rows; r += 2 * step) {
      const uint4 g0 = *reinterpret_cast<const uint4*>(dpost + r * N + c), h0 = *reinterpret_cast<const uint4*>(h + r * N + c);
      const uint4 g1 = *reinterpret_cast<const uint4*>(dpost + (r + step) * N + c), h1 = *reinterpret_cast<const uint4*>(h + (r + step) * N + c);
      one(g0, h0, r);
      one(g1, h1, r + step);
    }
    for (; r < rows; r += step)
      one(*reinterpret_cast<const uint4*>(dpost + r * N + c), *reinterpret_cast<const uint4*>(h + r * N + c), r);
  }
  if (colsum == nullptr) return;
#pragma unroll
  for (int j = 0; j < 8; ++j) s[ry][cp * 8 + j] = acc[j];
  __syncthreads();
  if (threadIdx.x < 128) {
    const int col = blockIdx.x * 128 + threadIdx.x;
    if (col < N) {
      float t = 0.f;
#pragma unroll
      for (int l = 0; l < 16; ++l) t += s[l][threadIdx.x];
      atomicAdd(&colsum[col], t);
    }
  }
}

extern "C" int iit_gelu_bwd_bias_colsum(const void* dpost, const void* h, const float* bias, void* dpre, float* colsum,
                                        int64_t rows, int64_t n, void* stream) {
  IIT_REQUIRE(dpost && h && dpre && rows >= 0 && n > 0 && n % 8 == 0 && n < (1ll << 31), "iit_gelu_bwd_bias_colsum: bad arguments "
              "(bf16 tensors, width a multiple of 8)");
  IIT_REQUIRE(aligned16(dpost, h, dpre, bias), "iit_gelu_bwd_bias_colsum: pointers must be 16-byte aligned");
  if (rows == 0) return 0;
  int64_t gy = (rows + 16 * 4 - 1) / (16 * 4);
  const int64_t gx = (n + 127) / 128, cap = std::max<int64_t>(1, (int64_t)kNumSMs * 8 / gx);
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  gelu_bwd_bias_colsum_kernel<<<dim3((unsigned)gx, (unsigned)gy), 256, 0, as_stream(stream)>>>(
      (const __nv_bfloat16*)dpost, (const __nv_bfloat16*)h, bias, (__nv_bfloat16*)dpre, colsum, rows, (int)n);
  IIT_LAUNCH_CHECK("iit_gelu_bwd_bias_colsum");
  return 0;
}

extern "C" int iit_colsum(const void* x, int dtype, float* out, int64_t rows, int64_t n, void* stream) {
  IIT_REQUIRE(x && out, "iit_colsum: null pointer");
  if (dtype == IIT_BF16 && n % 8 == 0 && n < (1ll << 31) && (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    int64_t gy = (rows + 16 * 8 - 1) / (16 * 8);          // >= 8 rows per thread
    const int64_t gx = (n + 127) / 128, cap = std::max<int64_t>(1, (int64_t)kNumSMs * 8 / gx);
    if (gy > cap) gy = cap;
    if (gy < 1) gy = 1;
    colsum_bf16x8_kernel<<<dim3((unsigned)gx, (unsigned)gy), 256, 0, as_stream(stream)>>>((const __nv_bfloat16*)x, out, rows, (int)n);
    IIT_LAUNCH_CHECK("iit_colsum");
    return 0;
  }
  int64_t by = (rows + 255) / 256;
  if (by > 64) by = 64;
  if (by < 1) by = 1;
  dim3 grid((unsigned)((n + 31) / 32), (unsigned)by);
  IIT_DISPATCH_DTYPE(dtype, T, { colsum_kernel<T><<<grid, 256, 0, as_stream(stream)>>>((const T*)x, out, rows, n); })
  IIT_LAUNCH_CHECK("iit_colsum");
  return 0;
}

extern "C" int iit_attn_fwd(const void* qkv, void* z, int dtype, int64_t batch, int seq, int heads, int d_head, void* stream) {
  IIT_REQUIRE(qkv && z && seq >= 1 && seq <= 32 && d_head >= 1 && d_head <= 32, "iit_attn_fwd: seq and d_head must be <= 32");
  const float scale = 1.0f / sqrtf((float)d_head);
  int64_t BH = batch * heads;
  if (dtype == IIT_BF16 && seq == 16 && d_head == 16 && aligned16(qkv, z)) {     // the IOI model: 16-byte loads and stores
    attn_fwd_s16_bf16_kernel<<<grid_for(BH, 8, 16), 128, 0, as_stream(stream)>>>((const __nv_bfloat16*)qkv, (__nv_bfloat16*)z, BH, heads, scale);
    IIT_LAUNCH_CHECK("iit_attn_fwd");
    return 0;
  }
  IIT_DISPATCH_DTYPE(dtype, T, {
    if (seq <= 16 && d_head <= 16)   // the IOI model: half the unrolled work, registers and shared memory of the general instance
      attn_fwd_kernel<T, 16, 16><<<grid_for(BH, 8, 16), 128, 0, as_stream(stream)>>>((const T*)qkv, (T*)z, BH, seq, heads, d_head, scale);
    else
      attn_fwd_kernel<T, 32, 32><<<grid_for(BH, 4), 128, 0, as_stream(stream)>>>((const T*)qkv, (T*)z, BH, seq, heads, d_head, scale);
  })
  IIT_LAUNCH_CHECK("iit_attn_fwd");
  return 0;
}
extern "C" int iit_attn_bwd(const void* qkv, const void* dz, void* dqkv, int dtype, int64_t batch, int seq, int heads, int d_head,
                            void* stream) {
  IIT_REQUIRE(qkv && dz && dqkv && seq >= 1 && seq <= 16 && d_head >= 1 && d_head <= 16,
              "iit_attn_bwd: seq and d_head must be <= 16");
  const float scale = 1.0f / sqrtf((float)d_head);
  int64_t BH = batch * heads;
  if (dtype == IIT_BF16 && seq == 16 && d_head == 16 && aligned16(qkv, dz, dqkv)) {
    attn_bwd_s16_bf16_kernel<<<grid_for(BH, 4, 32), 64, 0, as_stream(stream)>>>((const __nv_bfloat16*)qkv, (const __nv_bfloat16*)dz,
                                                                              (__nv_bfloat16*)dqkv, BH, heads, scale);
    IIT_LAUNCH_CHECK("iit_attn_bwd");
    return 0;
  }
  IIT_DISPATCH_DTYPE(dtype, T, {
    attn_bwd_kernel<T, 16, 16><<<grid_for(BH, 4, 32), 64, 0, as_stream(stream)>>>((const T*)qkv, (const T*)dz, (T*)dqkv, BH, seq, heads,
                                                                               d_head, scale);
  })
  IIT_LAUNCH_CHECK("iit_attn_bwd");
  return 0;
}
