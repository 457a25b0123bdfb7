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

template <typename T>
struct RasterBwdParams {
  FeatSet<const T> grad_out;   // n tensors (B,H,W,C_t)
  const int64_t* face_idx; // (B,H,W)
  const T* weights;        // (B,H,W,3) or nullptr => recompute
  const T* img;            // (B,F,3,2) NOT multiplied
  FeatSet<const T> feat;   // n tensors (B,F,3,C_t)
  T coord_mult;
  float pix_mult;
  float eps;
  int F;
  T* grad_img;             // (B,F,3,2)
  FeatSet<T> grad_feat;    // n tensors (B,F,3,C_t)
};

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

template <typename T>
__global__ void __launch_bounds__(256) raster_bwd_kernel(RasterBwdParams<T> bp, Geometry geo) {
  const int tile = blockIdx.x % geo.tiles_per_view;
  const int b = blockIdx.x / geo.tiles_per_view;
  const int tx = tile % geo.tiles_x, ty = tile / geo.tiles_x;
  const int px = tx * TILE + (threadIdx.x & (TILE - 1));
  const int py = ty * TILE + (threadIdx.x / TILE);
  const bool in_image = px < geo.W && py < geo.H;
  const int64_t pix = ((int64_t)b * geo.H + py) * geo.W + px;

  long long f = -1;
  if (in_image) f = bp.face_idx[pix];
  if (f >= bp.F) f = -1;                             // never write out of bounds on a corrupt index
  const bool covered = f >= 0;
  const unsigned any = __ballot_sync(0xffffffffu, covered);
  if (any == 0u) return;

  // group lanes by face (uncovered lanes form their own group and contribute zeros)
  const unsigned peers = __match_any_sync(0xffffffffu, (int)f);
  int src[5];
  group_partners(peers, src);
  unsigned steps = 0;
#pragma unroll
  for (int s = 0; s < 5; ++s)
    if (__any_sync(0xffffffffu, covered && src[s] != (int)(threadIdx.x & 31))) steps |= 1u << s;
  const bool leader = covered && ((peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u);

  T w0 = 0, w1 = 0, w2 = 0;
  T gv[6] = {0, 0, 0, 0, 0, 0};
  T acc1 = 0, acc2 = 0;
  const int64_t tf = (int64_t)b * bp.F + (covered ? f : 0);
  T k3 = 1, dw1[6], dw2[6];
  if (covered) {
    const T* ps = bp.img + tf * 6;
    const T ax = ps[0], ay = ps[1], bx = ps[2], by = ps[3], cx = ps[4], cy = ps[5];
    if (bp.weights) {
      const T* w = bp.weights + pix * 3;
      w0 = w[0]; w1 = w[1]; w2 = w[2];
    } else {
      typedef RN<T> R;
      const T x0 = (T)pix_coord(pix_scale(bp.pix_mult, geo.W), geo.W, px);
      const T y0 = (T)pix_coord(pix_scale(bp.pix_mult, geo.H), geo.H, geo.H - 1 - py);
      bary_weights<T>(R::mul(ax, bp.coord_mult), R::mul(ay, bp.coord_mult), R::mul(bx, bp.coord_mult),
                      R::mul(by, bp.coord_mult), R::mul(cx, bp.coord_mult), R::mul(cy, bp.coord_mult),
                      x0, y0, bp.eps, w0, w1, w2);
    }
    // rasterization_cuda.cu:302-374
    const T x0r = w0 * ax + w1 * bx + w2 * cx;
    const T y0r = w0 * ay + w1 * by + w2 * cy;
    const T m = bx - ax, p = by - ay, n = cx - ax, q = cy - ay, s = x0r - ax, t = y0r - ay;
    const T k1 = s * q - n * t;
    const T k2 = m * t - s * p;
    k3 = m * q - n * p;
    k3 = RN<T>::add_eps(k3, bp.eps);
    const T dw1dm = -q * k1, dw1dn = -t * k3 + p * k1, dw1dp = n * k1, dw1dq = s * k3 - m * k1;
    const T dw1ds = q * k3, dw1dt = -n * k3;
    const T dw2dm = t * k3 - q * k2, dw2dn = p * k2, dw2dp = -s * k3 + n * k2, dw2dq = -m * k2;
    const T dw2ds = -p * k3, dw2dt = m * k3;
    dw1[0] = -(dw1dm + dw1dn + dw1ds); dw1[1] = -(dw1dp + dw1dq + dw1dt);
    dw1[2] = dw1dm; dw1[3] = dw1dp; dw1[4] = dw1dn; dw1[5] = dw1dq;
    dw2[0] = -(dw2dm + dw2dn + dw2ds); dw2[1] = -(dw2dp + dw2dq + dw2dt);
    dw2[2] = dw2dm; dw2[3] = dw2dp; dw2[4] = dw2dn; dw2[5] = dw2dq;
  }
  const T inv_k3sq = (T)1 / (k3 * k3);               // one reciprocal instead of a division per channel (:389;
                                                     // gradients are judged at 1e-4 relative)
  for (int ft = 0; ft < bp.feat.n; ++ft) {
  const int C = bp.feat.ch[ft];
  const T* g = bp.grad_out.ptr[ft] + pix * C;
  const T* ff = bp.feat.ptr[ft] + tf * 3 * C;
  T* gf = bp.grad_feat.ptr[ft] + tf * 3 * C;
  for (int d = 0; d < C; ++d) {                     // C is warp-uniform
    T gd = 0;
    if (covered) {
      gd = g[d];
      const T c0 = ff[d], c1 = ff[C + d], c2 = ff[2 * C + d];
      // :382-389: dL/dI_d * ((c1-c0) * dw1[k] + (c2-c0) * dw2[k]) summed over the channels = acc1 * dw1[k] +
      // acc2 * dw2[k]: the six products are taken once per pixel, after the channel loop
      acc1 += gd * (c1 - c0);
      acc2 += gd * (c2 - c0);
    }
    // :273-285 feature gradients, aggregated per face inside the warp
    const T s0 = group_sum<T>(gd * w0, src, steps);
    const T s1 = group_sum<T>(gd * w1, src, steps);
    const T s2 = group_sum<T>(gd * w2, src, steps);
    if (leader) {
      atomicAdd(gf + d, s0);
      atomicAdd(gf + C + d, s1);
      atomicAdd(gf + 2 * C + d, s2);
    }
  }
  }
  T* gi = bp.grad_img + tf * 6;
  acc1 *= inv_k3sq; acc2 *= inv_k3sq;
  if (covered) {                                       // dw1 / dw2 are only defined on covered lanes
#pragma unroll
    for (int k = 0; k < 6; ++k) gv[k] = acc1 * dw1[k] + acc2 * dw2[k];
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    const T sk = group_sum<T>(gv[k], src, steps);
    if (leader) atomicAdd(gi + k, sk);               // :391-398
  }
}

}  // namespace dibr
