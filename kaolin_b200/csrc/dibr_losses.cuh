// dibr_losses.cuh — consumer of soft_mask: mask_iou of kaolin/metrics/render.py:18-40
//   up_b = sum(l*r), down_b = sum(l + r - l*r), loss = 1 - mean_b(up_b / (down_b + 1e-10))
// as a deterministic two-stage reduction (per-view chunk partials in fixed order, no float atomics) and an
// elementwise backward that yields grad_soft_mask analytically:
//   dloss/dl = -(1/B) * (r * (down+eps) - up * (1 - r)) / (down+eps)^2 * grad_loss
// The reference runs 8 torch kernels forward (mul, add, sub, 2 sums, add, div, mean) and their autograd
// backward, re-reading the (B,H,W) masks several times; here both masks are read once per direction.
#pragma once
#include "dibr_common.cuh"

namespace dibr {

constexpr int IOU_CHUNK = 8192;        // pixels per CTA (256 threads x 32)

template <typename T>
__device__ __forceinline__ T block_sum_256(T v, T* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) sh[w] = v;
  __syncthreads();
  T r = (T)0;
  if (w == 0) {
    r = l < 8 ? sh[l] : (T)0;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) r += __shfl_down_sync(0xffffffffu, r, o);
  }
  __syncthreads();
  return r;                            // valid in thread 0
}

// grid (chunks, B): partial[b][chunk] = {sum l*r, sum l+r-l*r}
template <typename T>
__global__ void __launch_bounds__(256) mask_iou_partial_kernel(const T* __restrict__ lhs, const T* __restrict__ rhs,
                                                               const int64_t P, const int chunks, T* partial) {
  __shared__ T sh[8];
  const int b = blockIdx.y;
  const int64_t base = (int64_t)b * P, lo = (int64_t)blockIdx.x * IOU_CHUNK;
  const int64_t hi = lo + IOU_CHUNK < P ? lo + IOU_CHUNK : P;
  T up = (T)0, dn = (T)0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += 256) {
    const T l = lhs[base + i], r = rhs[base + i];
    const T m = l * r;
    up += m;
    dn += (l + r) - m;
  }
  up = block_sum_256<T>(up, sh);
  dn = block_sum_256<T>(dn, sh);
  if (threadIdx.x == 0) {
    T* o = partial + ((int64_t)b * chunks + blockIdx.x) * 2;
    o[0] = up; o[1] = dn;
  }
}

// one CTA: per-view sums in fixed order -> sums (B,2) (saved for backward) and the scalar loss
template <typename T>
__global__ void __launch_bounds__(256) mask_iou_finish_kernel(const T* partial, const int B, const int chunks,
                                                              T* sums, T* loss) {
  __shared__ T sh[8];
  T acc = (T)0;
  for (int b = threadIdx.x; b < B; b += 256) {
    T up = (T)0, dn = (T)0;
    const T* p = partial + (int64_t)b * chunks * 2;
    for (int c = 0; c < chunks; ++c) { up += p[c * 2]; dn += p[c * 2 + 1]; }
    sums[b * 2] = up; sums[b * 2 + 1] = dn;
    acc += up / (dn + (T)1e-10);
  }
  acc = block_sum_256<T>(acc, sh);
  if (threadIdx.x == 0) loss[0] = (T)1.0 - acc / (T)B;
}

// grid (chunks, B).  grad_loss: device scalar (no host sync).  Either output may be nullptr.
template <typename T>
__global__ void __launch_bounds__(256) mask_iou_bwd_kernel(const T* __restrict__ lhs, const T* __restrict__ rhs,
                                                           const T* __restrict__ sums, const T* grad_loss,
                                                           const int64_t P, const int B, T* g_lhs, T* g_rhs) {
  const int b = blockIdx.y;
  const T up = sums[b * 2], de = sums[b * 2 + 1] + (T)1e-10;
  const T k = -grad_loss[0] / (T)B;
  const T a = k / de, c = k * up / (de * de);      // d/dx (up/de) = up'/de - up*de'/de^2
  const int64_t base = (int64_t)b * P, lo = (int64_t)blockIdx.x * IOU_CHUNK;
  const int64_t hi = lo + IOU_CHUNK < P ? lo + IOU_CHUNK : P;
  for (int64_t i = lo + threadIdx.x; i < hi; i += 256) {
    const T l = lhs[base + i], r = rhs[base + i];
    if (g_lhs) g_lhs[base + i] = a * r - c * ((T)1 - r);
    if (g_rhs) g_rhs[base + i] = a * l - c * ((T)1 - l);
  }
}

// ---------------------------------------------------------------- image L1 loss
// torch.mean(torch.abs(image - gt_image)) of the fitting loop (examples/tutorial/ian_dibr.py:260): the reference runs
// sub, abs and a mean reduction forward and sign / mul / div kernels backward, re-reading the images each time; here
// one deterministic two-stage reduction (fixed chunk order, no float atomics) and one elementwise backward
//   d loss / d a_i = sign(a_i - b_i) * grad_loss / n        (sign(0) = 0, as torch.abs' backward)
constexpr int L1_CHUNK = 16384;        // elements per CTA (256 threads x 64)

template <typename T>
__global__ void __launch_bounds__(256) image_l1_partial_kernel(const T* __restrict__ a, const T* __restrict__ b,
                                                               const int64_t n, T* partial) {
  __shared__ T sh[8];
  const int64_t lo = (int64_t)blockIdx.x * L1_CHUNK;
  const int64_t hi = lo + L1_CHUNK < n ? lo + L1_CHUNK : n;
  T acc = (T)0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += 256) {
    const T d = a[i] - b[i];
    acc += d < (T)0 ? -d : d;
  }
  acc = block_sum_256<T>(acc, sh);
  if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}
template <typename T>
__global__ void __launch_bounds__(256) image_l1_finish_kernel(const T* partial, const int chunks, const int64_t n,
                                                              T* loss) {
  __shared__ T sh[8];
  T acc = (T)0;
  for (int c = threadIdx.x; c < chunks; c += 256) acc += partial[c];
  acc = block_sum_256<T>(acc, sh);
  if (threadIdx.x == 0) loss[0] = acc / (T)n;          // n == 0: 0 / 0 = NaN, torch.mean of an empty tensor
}
template <typename T>
__global__ void __launch_bounds__(256) image_l1_bwd_kernel(const T* __restrict__ a, const T* __restrict__ b,
                                                           const T* grad_loss, const int64_t n, T* g_a, T* g_b) {
  const T k = grad_loss[0] / (T)n;
  const int64_t lo = (int64_t)blockIdx.x * L1_CHUNK;
  const int64_t hi = lo + L1_CHUNK < n ? lo + L1_CHUNK : n;
  for (int64_t i = lo + threadIdx.x; i < hi; i += 256) {
    const T d = a[i] - b[i];
    const T g = d > (T)0 ? k : (d < (T)0 ? -k : (d == d ? (T)0 : d));   // NaN propagates
    if (g_a) g_a[i] = g;
    if (g_b) g_b[i] = -g;
  }
}

// ---------------------------------------------------------------- Adam step of the fitting loop
// torch.optim.Adam (no weight decay, no amsgrad; examples/tutorial/ian_dibr.py:116-123 `vertices_optim`) for ONE
// parameter tensor, thread = element.  torch's fused / foreach implementations hand a tensor to `multi_tensor_apply`
// in chunks of 65 536 elements per CTA: the 60 k coordinates of a 20 k-vertex mesh are one CTA, 49 us of a 600 us
// step (profiles/r02_c4_b16_launches.csv).  `step` is a DEVICE counter (float, as torch's capturable Adam keeps it),
// incremented by the kernel's first thread after every thread has read it — no host value changes between replays
// of a captured step.
template <typename T>
__global__ void __launch_bounds__(256) adam_step_kernel(T* __restrict__ param, const T* __restrict__ grad,
                                                        T* __restrict__ exp_avg, T* __restrict__ exp_avg_sq,
                                                        float* step, unsigned* arrivals, const int64_t n,
                                                        const double lr, const double b1, const double b2,
                                                        const double eps) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const float t = step[0] + 1.0f;
  if (i < n) {
    const double g = (double)grad[i];
    const double m = b1 * (double)exp_avg[i] + (1.0 - b1) * g;
    const double v = b2 * (double)exp_avg_sq[i] + (1.0 - b2) * g * g;
    exp_avg[i] = (T)m;
    exp_avg_sq[i] = (T)v;
    const double bias1 = 1.0 - pow(b1, (double)t), bias2 = 1.0 - pow(b2, (double)t);
    const double denom = sqrt(v) / sqrt(bias2) + eps;
    param[i] = (T)((double)param[i] - (lr / bias1) * (m / denom));
  }
  // the last CTA to finish advances the counter (every CTA has read it by then)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(arrivals, 1u) == gridDim.x - 1) { step[0] = t; arrivals[0] = 0u; }
  }
}

}  // namespace dibr
