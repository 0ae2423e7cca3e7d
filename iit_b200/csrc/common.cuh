// Shared host/device helpers for the iit_b200 kernels (sm_100a).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/iit_b200.h"

namespace iit {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

void set_error(const char* fmt, ...);
void count_launch(int n = 1);
// which tensor-core kernel instantiation a dispatcher picked (diagnostic counters behind iit_kernel_variant_count)
enum KernelVariant {
  KV_TC_64 = 0, KV_TC_128, KV_TC_128_MT2, KV_TC_256, KV_TC2_256, KV_TC2_128, KV_WIN_3X3, KV_WIN_STEM, KV_WGRAD_TC,
  KV_WGRAD_WIN, KV_WGRAD_WIN_STEM, KV_COUNT
};
void count_variant(int id);

#define IIT_REQUIRE(cond, ...)            \
  do {                                    \
    if (!(cond)) {                        \
      iit::set_error(__VA_ARGS__);        \
      return 1;                           \
    }                                     \
  } while (0)

// Checked after every launch: configuration errors surface here, no sync is done.
#define IIT_LAUNCH_CHECK(name)                                                        \
  do {                                                                                \
    cudaError_t e__ = cudaGetLastError();                                             \
    if (e__ != cudaSuccess) {                                                         \
      iit::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__));         \
      return 2;                                                                       \
    }                                                                                 \
    iit::count_launch();                                                              \
  } while (0)

#define IIT_CUDA(call)                                                                \
  do {                                                                                \
    cudaError_t e__ = (call);                                                         \
    if (e__ != cudaSuccess) {                                                         \
      iit::set_error("%s failed: %s", #call, cudaGetErrorString(e__));                \
      return 3;                                                                       \
    }                                                                                 \
  } while (0)

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// Function attributes (opt-in dynamic shared memory) and occupancy answers are per DEVICE: one-time set-up is keyed
// by the current device ordinal, so a process that drives several GPUs sets every one of them up.
constexpr int kMaxDevices = 64;
inline int current_device_slot() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || d < 0) d = 0;
  return d % kMaxDevices;
}

inline int grid_for(int64_t work_items, int per_block, int max_waves = 8) {
  int64_t b = (work_items + per_block - 1) / per_block;
  int64_t cap = (int64_t)kNumSMs * max_waves;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// ---- element access -------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// 4 consecutive elements as floats (16 B for f32, 8 B for bf16); pointer must be so aligned.
template <typename T> struct Vec4;
template <> struct Vec4<float> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[4]) {
    float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct Vec4<__nv_bfloat16> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[4]) {
    uint2 t = *reinterpret_cast<const uint2*>(p);
    __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&t.x);
    __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&t.y);
    v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b);
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[4]) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]);
    __nv_bfloat162 b = __floats2bfloat162_rn(v[2], v[3]);
    uint2 t;
    t.x = *reinterpret_cast<uint32_t*>(&a);
    t.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = t;
  }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

#define IIT_DISPATCH_DTYPE(dtype, T, ...)                          \
  if ((dtype) == IIT_F32) { using T = float; __VA_ARGS__ }         \
  else if ((dtype) == IIT_BF16) { using T = __nv_bfloat16; __VA_ARGS__ } \
  else { iit::set_error("unsupported dtype %d", (int)(dtype)); return 1; }

}  // namespace iit
