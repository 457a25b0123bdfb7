// dibr_deftet.cuh — DefTet sparse (volumetric) renderer on the path's binning infrastructure
// (reference: kaolin/render/mesh/deftet.py:269-417, kaolin/csrc/render/mesh/deftet_cuda.cu).
//
// Semantics (deftet_cuda.cu:31-190 + the torch glue of DeftetSparseRenderer.forward, deftet.py:291-312): for
// every query point (pixel_coords are arbitrary image-plane points, not a grid) the FIRST knum faces in face
// order whose half-open bbox contains the point, whose barycentric weights are all >= 0 and whose depth lies
// in [min_depth, max_depth); those are then sorted by depth (descending) and their features interpolated.
//
// The reference scans all F faces per point (O(P*F), one warp per point).  Here:
//   D0  deftet_setup_count : thread = face.  bbox -> cell range on a uniform G x G grid over [-1,1]^2 (points
//       and bboxes outside are clamped to the border cells: the map is monotone, so a face is always listed in
//       the cell of every point its bbox contains) -> the SAME two-level lists as the rasterizer (fine cell /
//       4x4 coarse cell / per-view large list; scan_kernel and fill_kernel are reused unchanged).
//   D1  deftet_forward     : warp = query point.  Walks the three lists of the point's cell 32 faces at a time,
//       exact test with the reference's arithmetic (contractions read from its sm_100a SASS), ballot-compacted
//       append of (depth, face) keys into a per-warp shared-memory buffer; when the buffer would overflow it
//       is sorted by face and truncated to knum (which also yields an upper bound on the face index for the
//       rest of the walk); finally: first knum by face, bitonic sort by (depth desc, face asc), weights
//       recomputed from the face index (bit-identical), features interpolated, all knum slots written.
//   D2  deftet_backward    : thread = (point, k) entry; the reference's closed form (deftet_cuda.cu:290-400, the
//       same as the rasterizer's), red.global.add into the two gradient tensors.
#pragma once
#include "dibr_binning.cuh"

namespace dibr {

constexpr int DEFTET_WARPS = 4;                 // query points per CTA

template <typename T>
struct DeftetParams {
  const T* z;            // (B,F,3)
  const T* img;          // (B,F,3,2)
  const T* bbox;         // (B,F,4) xmin,ymin,xmax,ymax as given to the native op, or nullptr => min/max of img
  const T* pix;          // (B,P,2)
  const T* ranges;       // (B,P,2)  [min_depth, max_depth)
  int B, F, P, knum, grid; // grid = G cells per axis
  float eps;
  int cap;               // per-warp key buffer capacity (power of two, >= knum + 32)
  // reference-shaped outputs (insertion = face order), any may be nullptr
  int64_t* face_idx_unsorted; T* depth; T* w0; T* w1;
  // fused outputs (depth order)
  int64_t* face_idx;     // (B,P,knum)
  T* weights;            // (B,P,knum,3) or nullptr
  FeatSet<const T> feat; // inputs  (B,F,3,C_t)
  FeatSet<T> out;        // outputs (B,P,knum,C_t)
};

// monotone map image coordinate -> cell index (NaN -> 0; never matters: the exact test rejects NaN)
template <typename T>
__device__ __forceinline__ int deftet_cell(const T x, const int G) {
  const float c = ((float)x + 1.0f) * (0.5f * (float)G);
  if (!(c > 0.0f)) return 0;
  if (c >= (float)G) return G - 1;
  return (int)c;
}

template <typename T>
__global__ void __launch_bounds__(256) deftet_setup_count_kernel(const T* __restrict__ img,
                                                                 const T* __restrict__ bbox, FaceLayout fl,
                                                                 Geometry geo, BinLists L, const int G) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= fl.Ntot) return;
  T xmin, ymin, xmax, ymax;
  if (bbox) {
    xmin = bbox[g * 4]; ymin = bbox[g * 4 + 1]; xmax = bbox[g * 4 + 2]; ymax = bbox[g * 4 + 3];
  } else {                                  // deftet.py:288-290: torch.min / torch.max over the three vertices
    const T* v = img + g * 6;
    xmin = min_nan(min_nan(v[0], v[2]), v[4]); xmax = max_nan(max_nan(v[0], v[2]), v[4]);
    ymin = min_nan(min_nan(v[1], v[3]), v[5]); ymax = max_nan(max_nan(v[1], v[3]), v[5]);
  }
  uint2 rec = empty_range();
  // half-open bbox [min, max): empty (or NaN) boxes can never contain a point
  if (xmin < xmax && ymin < ymax) {
    // (float) of a double is monotone as well, so the cell range stays conservative in fp64
    const int cx0 = deftet_cell<T>(xmin, G), cx1 = deftet_cell<T>(xmax, G);
    const int cy0 = deftet_cell<T>(ymin, G), cy1 = deftet_cell<T>(ymax, G);
    // "virtual pixels": cell c spans pixels [c*TILE, c*TILE + TILE - 1], so that fill_kernel / tile_lists apply
    rec = pack_range(cx0 * TILE, cx1 * TILE + TILE - 1, cy0 * TILE, cy1 * TILE + TILE - 1);
    const int b = fl.view_of(g);
    int* cnt = L.tile_count + (int64_t)b * geo.slots_per_view;
    if ((cx1 - cx0 + 1) * (cy1 - cy0 + 1) <= PAIRS_PER_FACE) {
      for (int cy = cy0; cy <= cy1; ++cy)
        for (int cx = cx0; cx <= cx1; ++cx) atomicAdd(cnt + cy * geo.tiles_x + cx, 1);
    } else {
      const int qx0 = cx0 / COARSE, qx1 = cx1 / COARSE, qy0 = cy0 / COARSE, qy1 = cy1 / COARSE;
      if ((qx1 - qx0 + 1) * (qy1 - qy0 + 1) <= PAIRS_PER_FACE) {
        cnt += geo.tiles_per_view;
        for (int cy = qy0; cy <= qy1; ++cy)
          for (int cx = qx0; cx <= qx1; ++cx) atomicAdd(cnt + cy * geo.coarse_x + cx, 1);
      } else {
        const int64_t vb = fl.base(b);
        const int pos = atomicAdd(L.large_count + b, 1);
        L.large_list[vb + pos] = (int)(g - vb);
      }
    }
  }
  L.ranges[g] = rec;
}

// ---- exact per-(point, face) evaluation: deftet_cuda.cu:117-163, contractions as in the reference's SASS
//      (w = fma(p, q, -(r*s)); norm = (w0+w1)+w2; depth = fma(w2, cz, fma(w0, az, w1*bz))).
template <typename T>
__device__ __forceinline__ bool deftet_eval(const T* __restrict__ img6, const T* __restrict__ z3,
                                            const T* __restrict__ bbox4, const T x0,
                                            const T y0, const T min_depth, const T max_depth, const float eps,
                                            T& w0, T& w1, T& depth) {
  const T ax = img6[0], ay = img6[1], bx = img6[2], by = img6[3], cx = img6[4], cy = img6[5];
  T xmin, ymin, xmax, ymax;
  if (bbox4) {
    xmin = bbox4[0]; ymin = bbox4[1]; xmax = bbox4[2]; ymax = bbox4[3];
  } else {
    xmin = min_nan(min_nan(ax, bx), cx); xmax = max_nan(max_nan(ax, bx), cx);
    ymin = min_nan(min_nan(ay, by), cy); ymax = max_nan(max_nan(ay, by), cy);
  }
  if (!(x0 >= xmin && x0 < xmax && y0 >= ymin && y0 < ymax)) return false;
  const T aex = RN<T>::sub(ax, x0), aey = RN<T>::sub(ay, y0), bex = RN<T>::sub(bx, x0);
  const T bey = RN<T>::sub(by, y0), cex = RN<T>::sub(cx, x0), cey = RN<T>::sub(cy, y0);
  const T n0 = RN<T>::fma(bex, cey, -RN<T>::mul(bey, cex));
  const T n1 = RN<T>::fma(cex, aey, -RN<T>::mul(cey, aex));
  const T n2 = RN<T>::fma(aex, bey, -RN<T>::mul(aey, bex));
  const T norm = RN<T>::add(RN<T>::add(n0, n1), n2);
  const T den = RN<T>::add(norm, (T)copysignf(eps, (float)norm));
  w0 = RN<T>::div(n0, den);
  w1 = RN<T>::div(n1, den);
  const T w2 = RN<T>::div(n2, den);
  if (!(w0 >= (T)0 && w1 >= (T)0 && w2 >= (T)0)) return false;
  depth = RN<T>::fma(w2, z3[2], RN<T>::fma(w0, z3[0], RN<T>::mul(w1, z3[1])));
  return depth < max_depth && depth >= min_depth;
}

// order-preserving integer images of a depth: ascending key order == DESCENDING depth
__device__ __forceinline__ unsigned depth_key(float d) {
  const unsigned u = __float_as_uint(d);
  return ~(u ^ ((u & 0x80000000u) ? 0xffffffffu : 0x80000000u));
}
__device__ __forceinline__ unsigned long long depth_key(double d) {
  const unsigned long long u = (unsigned long long)__double_as_longlong(d);
  return ~(u ^ ((u >> 63) ? ~0ull : 0x8000000000000000ull));
}
template <typename T> struct DepthKey;
template <> struct DepthKey<float> { using type = unsigned; };
template <> struct DepthKey<double> { using type = unsigned long long; };

// warp-wide bitonic sort of n (<= N, N a power of two) (kd, kf) pairs in shared memory; entries >= n are
// padding and sort last.  by_face: ascending face index; otherwise ascending (depth key, face).
template <typename K>
__device__ __forceinline__ void warp_bitonic(K* kd, unsigned* kf, const int n, const bool by_face, const int lane) {
  int N = 32;
  while (N < n) N <<= 1;
  for (int i = n + lane; i < N; i += 32) { kd[i] = ~(K)0; kf[i] = 0xffffffffu; }
  __syncwarp();
  for (int k = 2; k <= N; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = lane; i < N; i += 32) {
        const int p = i ^ j;
        if (p > i) {
          const K da = kd[i], db = kd[p];
          const unsigned fa = kf[i], fb = kf[p];
          const bool gt = by_face ? (fa > fb) : (da > db || (da == db && fa > fb));
          const bool up = (i & k) == 0;
          if (gt == up) { kd[i] = db; kd[p] = da; kf[i] = fb; kf[p] = fa; }
        }
      }
      __syncwarp();
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(DEFTET_WARPS * 32) deftet_forward_kernel(DeftetParams<T> dp, FaceLayout fl,
                                                                           Geometry geo, BinLists L) {
  using K = typename DepthKey<T>::type;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  K* kd = reinterpret_cast<K*>(smem_raw) + (size_t)warp * dp.cap;
  unsigned* kf = reinterpret_cast<unsigned*>(reinterpret_cast<K*>(smem_raw) + (size_t)DEFTET_WARPS * dp.cap) +
                 (size_t)warp * dp.cap;
  const int64_t m = (int64_t)blockIdx.x * DEFTET_WARPS + warp;        // query point
  if (m >= (int64_t)dp.B * dp.P) return;
  const int b = (int)(m / dp.P);
  const T x0 = dp.pix[m * 2], y0 = dp.pix[m * 2 + 1];
  const T min_depth = dp.ranges[m * 2], max_depth = dp.ranges[m * 2 + 1];
  const int64_t vbase = fl.base(b);
  const T* img = dp.img + vbase * 6;
  const T* zz = dp.z + vbase * 3;
  const T* bb = dp.bbox ? dp.bbox + vbase * 4 : nullptr;
  const TileLists tl = tile_lists(L, geo, fl, b, deftet_cell<T>(x0, dp.grid), deftet_cell<T>(y0, dp.grid));
  const int total = tl.n_fine + tl.n_filtered();
  const int knum = dp.knum;
  int n = 0;
  unsigned thr = 0xffffffffu;                    // faces above thr cannot be among the first knum any more
  for (int c0 = 0; c0 < total; c0 += 32) {
    const int c = c0 + lane;
    unsigned f = 0xffffffffu;
    bool hit = false;
    T w0, w1, depth = (T)0;
    if (c < total) {
      f = (unsigned)(c < tl.n_fine ? tl.fine[c] : tl.filtered_at(c - tl.n_fine));
      if (f <= thr) hit = deftet_eval<T>(img + (int64_t)f * 6, zz + (int64_t)f * 3, bb ? bb + (int64_t)f * 4 : nullptr,
                                         x0, y0, min_depth, max_depth, dp.eps, w0, w1, depth);
    }
    unsigned bal = __ballot_sync(0xffffffffu, hit);
    if (bal == 0u) continue;
    if (n + __popc(bal) > dp.cap) {              // make room: keep the knum smallest faces seen so far
      warp_bitonic<K>(kd, kf, n, true, lane);
      if (n >= knum) { n = knum; thr = kf[knum - 1]; }
      __syncwarp();
      hit = hit && f <= thr;
      bal = __ballot_sync(0xffffffffu, hit);
    }
    if (hit) {
      const int pos = n + __popc(bal & ((1u << lane) - 1u));
      kd[pos] = depth_key(depth);
      kf[pos] = f;
    }
    n += __popc(bal);
    __syncwarp();
  }
  if (n > knum) {
    warp_bitonic<K>(kd, kf, n, true, lane);
    n = knum;
  }
  const bool unsorted_out = dp.face_idx_unsorted != nullptr;
  if (n > 1) warp_bitonic<K>(kd, kf, n, unsorted_out, lane);
  __syncwarp();
  // emit all knum slots
  for (int e = lane; e < knum; e += 32) {
    const int64_t o = m * knum + e;
    int64_t fidx = -1;
    T w0 = (T)0, w1 = (T)0, depth = -INFINITY;
    if (e < n) {
      const unsigned f = kf[e];
      fidx = (int64_t)f;
      deftet_eval<T>(img + (int64_t)f * 6, zz + (int64_t)f * 3, bb ? bb + (int64_t)f * 4 : nullptr, x0, y0, min_depth,
                     max_depth, dp.eps, w0, w1, depth);
    }
    if (unsorted_out) {                           // native-op shape: deftet.cpp:87-92 fills + kernel writes
      dp.face_idx_unsorted[o] = fidx;
      dp.depth[o] = depth; dp.w0[o] = w0; dp.w1[o] = w1;
      continue;
    }
    dp.face_idx[o] = fidx;
    // deftet.py:296: w2 = (face_idx != -1).float() - (w0 + w1)
    const T w2 = (e < n ? (T)1 : (T)0) - (w0 + w1);
    if (dp.weights) { T* w = dp.weights + o * 3; w[0] = w0; w[1] = w1; w[2] = w2; }
    for (int t = 0; t < dp.out.n; ++t) {
      const int C = dp.out.ch[t];
      T* out = dp.out.ptr[t] + o * C;
      if (e < n) {
        const T* ff = dp.feat.ptr[t] + (vbase + (int64_t)fidx) * 3 * C;
        // deftet.py:309-310: sum over the three vertices of weight * feature, in vertex order
        for (int d = 0; d < C; ++d) out[d] = (w0 * ff[d] + w1 * ff[C + d]) + w2 * ff[2 * C + d];
      } else {
        for (int d = 0; d < C; ++d) out[d] = (T)0;
      }
    }
  }
}

// ---- backward: thread = (point, k) entry; deftet_cuda.cu:240-410
template <typename T>
struct DeftetBwdParams {
  const int64_t* face_idx;   // (B,P,knum)
  const T* weights;          // (B,P,knum,3)
  const T* img;              // (B,F,3,2)
  int B, F, P, knum;
  float eps;
  FeatSet<const T> grad_out; // (B,P,knum,C_t)
  FeatSet<const T> feat;     // (B,F,3,C_t)
  FeatSet<T> grad_feat;      // (B,F,3,C_t)  zero-filled by the host wrapper
  T* grad_img;               // (B,F,3,2)    zero-filled by the host wrapper
};

template <typename T>
__global__ void __launch_bounds__(256) deftet_backward_kernel(DeftetBwdParams<T> bp) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (int64_t)bp.B * bp.P * bp.knum) return;
  const int64_t f = bp.face_idx[e];
  if (f < 0 || f >= bp.F) return;
  const int b = (int)(e / ((int64_t)bp.P * bp.knum));
  const int64_t g = (int64_t)b * bp.F + f;
  const T aw = bp.weights[e * 3], bw = bp.weights[e * 3 + 1], cw = bp.weights[e * 3 + 2];
  const T* v = bp.img + g * 6;
  const T ax = v[0], ay = v[1], bx = v[2], by = v[3], cx = v[4], cy = v[5];
  const T x0 = aw * ax + bw * bx + cw * cx, y0 = aw * ay + bw * by + cw * cy;
  const T m = bx - ax, p = by - ay, n = cx - ax, q = cy - ay, s = x0 - ax, t = y0 - ay;
  const T k1 = s * q - n * t, k2 = m * t - s * p;
  T k3 = m * q - n * p;
  k3 += (T)copysign((double)bp.eps, (double)k3);
  const T inv = (T)1 / (k3 * k3);
  const T dw1dm = -q * k1, dw1dn = p * k1 - t * k3, dw1dp = n * k1, dw1dq = s * k3 - m * k1;
  const T dw1ds = q * k3, dw1dt = -n * k3;
  const T dw2dm = t * k3 - q * k2, dw2dn = p * k2, dw2dp = n * k2 - s * k3, dw2dq = -m * k2;
  const T dw2ds = -p * k3, dw2dt = m * k3;
  const T dw1dax = -(dw1dm + dw1dn + dw1ds), dw1day = -(dw1dp + dw1dq + dw1dt);
  const T dw2dax = -(dw2dm + dw2dn + dw2ds), dw2day = -(dw2dp + dw2dq + dw2dt);
  T acc1 = (T)0, acc2 = (T)0;                     // sum_c dL/dI_c * (c1-c0), sum_c dL/dI_c * (c2-c0)
  for (int tt = 0; tt < bp.feat.n; ++tt) {
    const int C = bp.feat.ch[tt];
    const T* go = bp.grad_out.ptr[tt] + e * C;
    const T* ff = bp.feat.ptr[tt] + g * 3 * C;
    T* gf = bp.grad_feat.ptr[tt] ? bp.grad_feat.ptr[tt] + g * 3 * C : nullptr;
    for (int d = 0; d < C; ++d) {
      const T gd = go[d];
      if (gf && gd != (T)0) {
        atomicAdd(gf + d, gd * aw);
        atomicAdd(gf + C + d, gd * bw);
        atomicAdd(gf + 2 * C + d, gd * cw);
      }
      const T c0 = ff[d];
      acc1 += gd * (ff[C + d] - c0);
      acc2 += gd * (ff[2 * C + d] - c0);
    }
  }
  if (!bp.grad_img) return;
  acc1 *= inv; acc2 *= inv;
  T* gi = bp.grad_img + g * 6;
  atomicAdd(gi + 0, acc1 * dw1dax + acc2 * dw2dax);
  atomicAdd(gi + 1, acc1 * dw1day + acc2 * dw2day);
  atomicAdd(gi + 2, acc1 * dw1dm + acc2 * dw2dm);
  atomicAdd(gi + 3, acc1 * dw1dp + acc2 * dw2dp);
  atomicAdd(gi + 4, acc1 * dw1dn + acc2 * dw2dn);
  atomicAdd(gi + 5, acc1 * dw1dq + acc2 * dw2dq);
}

}  // namespace dibr
