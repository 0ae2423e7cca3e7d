// Either side of the training step: building batches of (base, source) pairs and the
// sweep's counterfactual inputs on the device.
//   * iit_gather_rows:  dst[b, :] = src[idx[b], :]  -- the collate step of IITDataset (reference:
//     iit/utils/iit_dataset.py:24-59 stacks per-sample host tensors and copies each one to the device) as one
//     coalesced gather over a device-resident dataset; the pair indices themselves stay on the host, bit-exact.
//   * iit_gather_box:   dst[b, :, r0:r0+h, c0:c0+w] = bank[idx[b]]  -- the quadrant swap of
//     MNIST_PVR_Dataset.patch_at_hl_idx (reference: iit/tasks/mnist_pvr/dataset.py:125-155): every image of the batch gets
//     one quadrant replaced by a digit drawn from a device-resident bank.
//   * iit_pick_other_class: the "different class" draw of that function without its per-sample rejection loop: with the
//     bank sorted by label, a uniform draw over the items of every OTHER class is one integer per sample.
// HBM-bound byte movement: 16-byte units, consecutive lanes on consecutive units of one row.
#include "common.cuh"

namespace iit {

template <typename U>
__global__ void __launch_bounds__(256) gather_rows_kernel(U* __restrict__ dst, const U* __restrict__ src,
                                                          const int64_t* __restrict__ idx, int64_t rows, int64_t units,
                                                          int64_t src_rows, int* __restrict__ err) {
  const int64_t total = rows * units;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / units, u = i - b * units;
    const int64_t s = idx[b];
    if (s < 0 || s >= src_rows) { *err = 1; continue; }
    dst[i] = src[s * units + u];
  }
}

// one block per (image, channel); threads walk the h x w box row by row
template <typename T>
__global__ void __launch_bounds__(128) gather_box_kernel(T* __restrict__ dst, int64_t sn, int64_t sc, int64_t sh, int64_t sw,
                                                         const T* __restrict__ bank, const int64_t* __restrict__ idx, int C, int h,
                                                         int w, int r0, int c0, int64_t bank_items, int* __restrict__ err) {
  const int b = blockIdx.x / C, c = blockIdx.x - b * C;
  const int64_t s = idx[b];
  if (s < 0 || s >= bank_items) { if (threadIdx.x == 0) *err = 1; return; }
  const T* src = bank + (s * C + c) * (int64_t)h * w;
  T* d = dst + b * sn + c * sc + r0 * sh + c0 * sw;
  for (int i = threadIdx.x; i < h * w; i += blockDim.x) {
    const int y = i / w, x = i - y * w;
    d[y * sh + x * sw] = src[i];
  }
}

// splitmix64: counter-based, so sample b of draw `seed` does not depend on the launch geometry
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

__global__ void pick_other_class_kernel(const int64_t* __restrict__ cur, const int64_t* __restrict__ class_start, int n_classes,
                                        int64_t bank_items, uint64_t seed, int64_t n, int64_t* __restrict__ out_idx,
                                        int64_t* __restrict__ out_label, int* __restrict__ err) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n) return;
  const int64_t c = cur[b];
  if (c < 0 || c >= n_classes) { *err = 1; return; }
  const int64_t lo = class_start[c], hi = class_start[c + 1];       // the bank is sorted by label
  const int64_t others = bank_items - (hi - lo);
  if (others <= 0) { *err = 2; return; }
  // multiply-shift maps 64 random bits to [0, others) (bias < 2^-40 for banks of this size)
  const uint64_t r = splitmix64(seed * 0x100000001B3ull + (uint64_t)b);
  int64_t k = (int64_t)__umul64hi(r, (uint64_t)others);
  if (k >= lo) k += hi - lo;                                         // skip the current class's range
  out_idx[b] = k;
  // label of item k: the class whose range contains it (n_classes is small)
  int64_t lab = 0;
  for (int j = 0; j < n_classes; ++j) if (k >= class_start[j]) lab = j;
  out_label[b] = lab;
}

}  // namespace iit

using namespace iit;

extern "C" int iit_gather_rows(void* dst, const void* src, const int64_t* idx, int64_t rows, int64_t row_bytes, int64_t src_rows,
                               int* err_flag, void* stream) {
  IIT_REQUIRE(dst && src && idx && err_flag && rows >= 0 && row_bytes > 0 && src_rows > 0, "iit_gather_rows: bad arguments");
  if (rows == 0) return 0;
  const uintptr_t m = reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src) | (uintptr_t)row_bytes;
  cudaStream_t st = as_stream(stream);
#define IIT_GATHER(U)                                                                                              \
  gather_rows_kernel<U><<<grid_for(rows * (row_bytes / (int64_t)sizeof(U)), 256 * 4), 256, 0, st>>>(             \
      (U*)dst, (const U*)src, idx, rows, row_bytes / (int64_t)sizeof(U), src_rows, err_flag)
  if ((m & 15) == 0) IIT_GATHER(uint4);
  else if ((m & 7) == 0) IIT_GATHER(uint2);
  else if ((m & 3) == 0) IIT_GATHER(uint32_t);
  else if ((m & 1) == 0) IIT_GATHER(uint16_t);
  else IIT_GATHER(uint8_t);
#undef IIT_GATHER
  IIT_LAUNCH_CHECK("iit_gather_rows");
  return 0;
}

extern "C" int iit_gather_box(void* dst, const int64_t* dst_strides_nchw, const void* bank, const int64_t* idx, int dtype, int n,
                              int c, int h, int w, int r0, int c0, int64_t bank_items, int* err_flag, void* stream) {
  IIT_REQUIRE(dst && dst_strides_nchw && bank && idx && err_flag && c >= 1 && h >= 1 && w >= 1 && r0 >= 0 && c0 >= 0 && bank_items > 0,
              "iit_gather_box: bad arguments");
  if (n == 0) return 0;
  IIT_REQUIRE((int64_t)n * c < (1ll << 31), "iit_gather_box: too many (image, channel) planes");
  cudaStream_t st = as_stream(stream);
  IIT_DISPATCH_DTYPE(dtype, T, {
    gather_box_kernel<T><<<(unsigned)(n * c), 128, 0, st>>>((T*)dst, dst_strides_nchw[0], dst_strides_nchw[1], dst_strides_nchw[2],
                                                           dst_strides_nchw[3], (const T*)bank, idx, c, h, w, r0, c0, bank_items,
                                                           err_flag);
  })
  IIT_LAUNCH_CHECK("iit_gather_box");
  return 0;
}

extern "C" int iit_pick_other_class(const int64_t* cur_class, const int64_t* class_start, int n_classes, int64_t bank_items,
                                    uint64_t seed, int64_t n, int64_t* out_idx, int64_t* out_label, int* err_flag, void* stream) {
  IIT_REQUIRE(cur_class && class_start && out_idx && out_label && err_flag && n_classes >= 2 && bank_items > 0,
              "iit_pick_other_class: bad arguments");
  if (n == 0) return 0;
  pick_other_class_kernel<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(cur_class, class_start, n_classes, bank_items,
                                                                                    seed, n, out_idx, out_label, err_flag);
  IIT_LAUNCH_CHECK("iit_pick_other_class");
  return 0;
}
