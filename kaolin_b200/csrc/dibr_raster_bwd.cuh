// dibr_raster_bwd.cuh — K4: gradients of the interpolated features w.r.t. the image-plane vertex
// positions and the per-vertex face features (reference: rasterization_cuda.cu:238-402).
//
// One thread per pixel, laid out tile-wise (a warp covers a 16x2 pixel block) so that the lanes of a
// warp mostly hit the same few faces.  Lanes are grouped by face with __match_any_sync, each group
// tree-reduces its 6 + 3C partial sums by rank with shuffles, and only the group leader issues
// red.global.add — one atomic per (warp, face, component) instead of 9C per pixel.
// The barycentric weights are either read (reference-shaped call) or recomputed bit-exactly with the
// forward arithmetic, which removes the (B,H,W,3) save/reload.
#pragma once
#include "dibr_binning.cuh"

namespace dibr {

#ifndef DIBR_RASTER_BWD_CTAS
#define DIBR_RASTER_BWD_CTAS 6     // resident 256-thread CTAs per SM the register allocation aims at.  The kernel waits
                                   // on its dependent gathers (face_idx -> coordinates -> features; ncu: long-scoreboard),
                                   // so occupancy beats registers — measured after the derivative terms moved behind the
                                   // channel loop (c3-64 / c5 / c2): 4 CTAs, 64 registers, no spills 1.275 / 0.667 /
                                   // 0.209 ms; 6 CTAs, 40 registers, 40 B of spills **1.093 / 0.588 / 0.209**; 8 CTAs, 32
                                   // registers, 88 B 1.121 / 0.618 / 0.227
#endif

#ifndef DIBR_BWD_PREFETCH
#define DIBR_BWD_PREFETCH 0
#endif
#ifndef DIBR_BWD_MAXSTEPS
#define DIBR_BWD_MAXSTEPS 5        // shuffle steps of the per-face reduction inside a warp (see raster_bwd_tile)
#endif

template <typename T>
struct RasterBwdParams {
  FeatSet<const T> grad_out;   // n tensors (B,H,W,C_t)
  const int64_t* face_idx; // (B,H,W)
  const T* weights;        // (B,H,W,3) or nullptr => recompute
  const T* img;            // (B,F,3,2) NOT multiplied
  FeatSet<const T> feat;   // n tensors (B,F,3,C_t)
  T coord_mult;
  float pix_mult;
  float sx, sy;            // pix_scale(pix_mult, W / H): one IEEE division each, taken on the host (same fp32 result)
  float eps;
  int F;
  T* grad_img;             // (B,F,3,2)
  FeatSet<T> grad_feat;    // n tensors (B,F,3,C_t)
};

// img gradient of one (warp, face): three 8-byte vector reductions (red.global.add.v2.f32, sm_90+) instead of six
// scalar ones — a face's (3,2) record is 24 bytes, so every vertex pair is 8-byte aligned
__device__ __forceinline__ void red_add2(float* p, const float a, const float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void red_add2(double* p, const double a, const double b) {
  atomicAdd(p, a);
  atomicAdd(p + 1, b);
}

// Sum `v` over the lanes of `peers` (every lane of the warp must call this with its own peers mask);
// the result is valid on the group's lowest lane.  `src[s]` = partner lane for step s (self if none): the member
// 2^s ranks further in the group, found by pointer jumping over "next member of my group" (4 shuffles).
__device__ __forceinline__ void group_partners(unsigned peers, int* src) {
  const int lane = threadIdx.x & 31;
  const int rank = __popc(peers & ((1u << lane) - 1u));
  const unsigned above = peers & ~(0xffffffffu >> (31 - lane));
  int nxt = above ? (__ffs(above) - 1) : -1;              // member with rank + 1
#pragma unroll
  for (int s = 0; s < 5; ++s) {
    const int step = 1 << s;
    src[s] = ((rank & (2 * step - 1)) == 0 && nxt >= 0) ? nxt : lane;
    if (s < 4) {
      const int hop = __shfl_sync(0xffffffffu, nxt, nxt >= 0 ? nxt : lane);   // nxt of my nxt: rank + 2 * step
      nxt = nxt >= 0 ? hop : -1;
    }
  }
}
template <typename T>
__device__ __forceinline__ T group_sum(T v, const int* src, unsigned active_steps) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int s = 0; s < 5; ++s) {
    if (active_steps & (1u << s)) {                 // warp-uniform: skip steps no group needs
      const T o = __shfl_sync(0xffffffffu, v, src[s]);
      if (src[s] != lane) v += o;
    }
  }
  return v;
}

// group_sum with a compile-time number of steps (steps 0..NS-1; a step no group needs adds nothing)
template <typename T, int NS>
__device__ __forceinline__ T group_sum_n(T v, const int* src) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int s = 0; s < NS; ++s) {
    const T o = __shfl_sync(0xffffffffu, v, src[s]);
    if (src[s] != lane) v += o;
  }
  return v;
}

// CC channels (d .. d+CC-1 of a tensor with C channels) of one pixel: loads batched, :273-285 feature gradients
// aggregated per face inside the warp, contributions to the coordinate gradient folded into acc1 / acc2.
template <typename T, int NS, int CC>
__device__ __forceinline__ void raster_bwd_block(const T* __restrict__ g, const T* __restrict__ ff, T* gf, const int C,
                                                 const bool covered, const bool leader, const T w0, const T w1,
                                                 const T w2, const int* src, T& acc1, T& acc2) {
  T gd[CC], c0[CC], c1[CC], c2[CC];
#pragma unroll
  for (int d = 0; d < CC; ++d) {
    gd[d] = 0; c0[d] = 0; c1[d] = 0; c2[d] = 0;
    if (covered) { gd[d] = g[d]; c0[d] = ff[d]; c1[d] = ff[C + d]; c2[d] = ff[2 * C + d]; }
  }
#pragma unroll
  for (int d = 0; d < CC; ++d) {
    // :382-389: dL/dI_d * ((c1-c0) * dw1[k] + (c2-c0) * dw2[k]) summed over the channels = acc1 * dw1[k] +
    // acc2 * dw2[k]: the six products are taken once per pixel, after the channel loop
    acc1 += gd[d] * (c1[d] - c0[d]);
    acc2 += gd[d] * (c2[d] - c0[d]);
    const T s0 = group_sum_n<T, NS>(gd[d] * w0, src);
    const T s1 = group_sum_n<T, NS>(gd[d] * w1, src);
    const T s2 = group_sum_n<T, NS>(gd[d] * w2, src);
    if (leader) {
      atomicAdd(gf + d, s0);
      atomicAdd(gf + C + d, s1);
      atomicAdd(gf + 2 * C + d, s2);
    }
  }
}

// Everything after the weights, for a warp whose largest group needs NS reduction steps: the feature gradients
// (channel loop), then the coordinate gradient.  The derivative terms dw1 / dw2 (rasterization_cuda.cu:302-374) are
// formed AFTER the channel loop from the face's coordinates, re-read from L1: held across the loop they pushed the
// kernel into a local-memory frame (40 LDL / 10 STL in the round-1 SASS even at 64 registers).
template <typename T, int NS>
__device__ __forceinline__ void raster_bwd_tail(const RasterBwdParams<T>& bp, const int64_t pix, const int64_t tf,
                                                const bool covered, const bool leader, const T w0, const T w1,
                                                const T w2, const int* src) {
  T acc1 = 0, acc2 = 0;
  for (int ft = 0; ft < bp.feat.n; ++ft) {
    const int C = bp.feat.ch[ft];                      // warp-uniform
    const T* g = bp.grad_out.ptr[ft] + pix * C;
    const T* ff = bp.feat.ptr[ft] + tf * 3 * C;
    T* gf = bp.grad_feat.ptr[ft] + tf * 3 * C;
    int d = 0;
    for (; d + 4 <= C; d += 4)
      raster_bwd_block<T, NS, 4>(g + d, ff + d, gf + d, C, covered, leader, w0, w1, w2, src, acc1, acc2);
    if (d + 2 <= C) {
      raster_bwd_block<T, NS, 2>(g + d, ff + d, gf + d, C, covered, leader, w0, w1, w2, src, acc1, acc2);
      d += 2;
    }
    if (d < C) raster_bwd_block<T, NS, 1>(g + d, ff + d, gf + d, C, covered, leader, w0, w1, w2, src, acc1, acc2);
  }
  T dw1[6], dw2[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) { dw1[k] = 0; dw2[k] = 0; }
  if (covered) {
    const T* ps = bp.img + tf * 6;
    const T ax = ps[0], ay = ps[1], bx = ps[2], by = ps[3], cx = ps[4], cy = ps[5];
    // rasterization_cuda.cu:302-374
    const T x0r = w0 * ax + w1 * bx + w2 * cx;
    const T y0r = w0 * ay + w1 * by + w2 * cy;
    const T m = bx - ax, p = by - ay, n = cx - ax, q = cy - ay, s = x0r - ax, t = y0r - ay;
    const T k1 = s * q - n * t;
    const T k2 = m * t - s * p;
    T k3 = m * q - n * p;
    k3 = RN<T>::add_eps(k3, bp.eps);
    const T inv_k3sq = (T)1 / (k3 * k3);             // one reciprocal instead of a division per channel (:389;
                                                     // gradients are judged at 1e-4 relative)
    acc1 *= inv_k3sq; acc2 *= inv_k3sq;
    const T dw1dm = -q * k1, dw1dn = -t * k3 + p * k1, dw1dp = n * k1, dw1dq = s * k3 - m * k1;
    const T dw1ds = q * k3, dw1dt = -n * k3;
    const T dw2dm = t * k3 - q * k2, dw2dn = p * k2, dw2dp = -s * k3 + n * k2, dw2dq = -m * k2;
    const T dw2ds = -p * k3, dw2dt = m * k3;
    dw1[0] = -(dw1dm + dw1dn + dw1ds); dw1[1] = -(dw1dp + dw1dq + dw1dt);
    dw1[2] = dw1dm; dw1[3] = dw1dp; dw1[4] = dw1dn; dw1[5] = dw1dq;
    dw2[0] = -(dw2dm + dw2dn + dw2ds); dw2[1] = -(dw2dp + dw2dq + dw2dt);
    dw2[2] = dw2dm; dw2[3] = dw2dp; dw2[4] = dw2dn; dw2[5] = dw2dq;
  }
  T* gi = bp.grad_img + tf * 6;
#pragma unroll
  for (int k = 0; k < 6; k += 2) {
    const T g0 = covered ? acc1 * dw1[k] + acc2 * dw2[k] : (T)0;
    const T g1 = covered ? acc1 * dw1[k + 1] + acc2 * dw2[k + 1] : (T)0;
    const T s0 = group_sum_n<T, NS>(g0, src);
    const T s1 = group_sum_n<T, NS>(g1, src);
    if (leader) red_add2(gi + k, s0, s1);            // :391-398
  }
}

template <typename T>
__device__ __forceinline__ void raster_bwd_tile(const RasterBwdParams<T>& bp, const Geometry& geo, const int b,
                                                const int tx, const int ty);

// grid (tiles_x, tiles_y, min(B, 65535)): no integer division in the per-thread prologue
template <typename T>
__global__ void __launch_bounds__(256, DIBR_RASTER_BWD_CTAS) raster_bwd_kernel(RasterBwdParams<T> bp, Geometry geo) {
  for (int b = blockIdx.z; b < geo.B; b += gridDim.z) raster_bwd_tile<T>(bp, geo, b, blockIdx.x, blockIdx.y);
}

template <typename T>
__device__ __forceinline__ void raster_bwd_tile(const RasterBwdParams<T>& bp, const Geometry& geo, const int b,
                                                const int tx, const int ty) {
  const int px = tx * TILE + (threadIdx.x & (TILE - 1));
  const int py = ty * TILE + (threadIdx.x / TILE);
  const bool in_image = px < geo.W && py < geo.H;
  const int64_t pix = ((int64_t)b * geo.H + py) * geo.W + px;

  long long f = -1;
  if (in_image) f = bp.face_idx[pix];
#if DIBR_BWD_PREFETCH
  // the upstream gradient's address depends on the pixel only: start its fetch together with face_idx, two dependent
  // gathers (face_idx -> coordinates / features) before it is used (ncu: the kernel waits on long-scoreboard)
  if (in_image) {
    for (int ft = 0; ft < bp.feat.n; ++ft)
      asm volatile("prefetch.global.L1 [%0];" ::"l"(bp.grad_out.ptr[ft] + pix * bp.feat.ch[ft]));
  }
#endif
  if (f >= bp.F) f = -1;                             // never write out of bounds on a corrupt index
  const bool covered = f >= 0;
  const unsigned any = __ballot_sync(0xffffffffu, covered);
  if (any == 0u) return;                             // (warp-uniform)

  // group lanes by face (uncovered lanes form their own group and contribute zeros)
  const unsigned peers = __match_any_sync(0xffffffffu, (int)f);
  int src[5];
  group_partners(peers, src);
  unsigned steps = 0;
#pragma unroll
  for (int s = 0; s < 5; ++s)
    if (__any_sync(0xffffffffu, covered && src[s] != (int)(threadIdx.x & 31))) steps |= 1u << s;
  // A group is reduced in sub-groups of 2^DIBR_BWD_MAXSTEPS consecutive ranks, each with its own leader (the member
  // whose rank is a multiple of that size): 5 = one leader per (warp, face).
  const int my_rank = __popc(peers & ((1u << (threadIdx.x & 31)) - 1u));
  const bool leader = covered && (my_rank & ((1 << DIBR_BWD_MAXSTEPS) - 1)) == 0;
  steps &= (1u << DIBR_BWD_MAXSTEPS) - 1u;
#pragma unroll
  for (int s = DIBR_BWD_MAXSTEPS; s < 5; ++s) src[s] = (int)(threadIdx.x & 31);     // steps beyond the sub-group: none

  T w0 = 0, w1 = 0, w2 = 0;
  const int64_t tf = (int64_t)b * bp.F + (covered ? f : 0);
  if (covered) {
    if (bp.weights) {
      const T* w = bp.weights + pix * 3;
      w0 = w[0]; w1 = w[1]; w2 = w[2];
    } else {
      typedef RN<T> R;
      const T* ps = bp.img + tf * 6;
      const T ax = ps[0], ay = ps[1], bx = ps[2], by = ps[3], cx = ps[4], cy = ps[5];
      const T x0 = (T)pix_coord(bp.sx, geo.W, px);
      const T y0 = (T)pix_coord(bp.sy, geo.H, geo.H - 1 - py);
      bary_weights<T>(R::mul(ax, bp.coord_mult), R::mul(ay, bp.coord_mult), R::mul(bx, bp.coord_mult),
                      R::mul(by, bp.coord_mult), R::mul(cx, bp.coord_mult), R::mul(cy, bp.coord_mult),
                      x0, y0, bp.eps, w0, w1, w2);
    }
  }
  // the reduction depth of the warp's largest group, dispatched once: no per-step branches in the channel loops
  const int ns = 32 - __clz(steps);
  if (DIBR_BWD_MAXSTEPS <= 2 || ns <= 2) raster_bwd_tail<T, (DIBR_BWD_MAXSTEPS < 2 ? DIBR_BWD_MAXSTEPS : 2)>(bp, pix, tf, covered, leader, w0, w1, w2, src);
  else if (ns <= 3) raster_bwd_tail<T, 3>(bp, pix, tf, covered, leader, w0, w1, w2, src);
  else if (ns == 4) raster_bwd_tail<T, 4>(bp, pix, tf, covered, leader, w0, w1, w2, src);
  else raster_bwd_tail<T, 5>(bp, pix, tf, covered, leader, w0, w1, w2, src);
}

}  // namespace dibr
