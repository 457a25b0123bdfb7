// dibr_raster_fwd.cuh — K2: per-tile depth-tested face search + feature interpolation.
//
// One CTA = one 16x16 pixel tile of one view, one thread = one pixel.  The tile's face list (from
// dibr_binning.cuh) plus the view's "large" faces are staged through shared memory in batches of 256
// (warp-ballot / prefix-scan compaction for the filtered large list); every pixel walks the batch with
// broadcast LDS reads, applies the integer pixel-range test (exactly equivalent to the reference's float
// bbox test, see count_below) and the reference's barycentric arithmetic.
//
// Winner rule (order independent, equals the reference's in-order scan rasterization_cuda.cu:88-171):
//   * among faces whose z0 is not NaN: the largest z0 (> -inf), lowest face index on ties;
//   * if any inside face has a NaN z0 the reference's `z0 <= max_z0` is false from then on, so the
//     LAST inside face in index order wins — tracked as (has_nan, max inside index).
#pragma once
#include "dibr_binning.cuh"

namespace dibr {

template <typename T>
struct RasterFwdParams {
  const T* z;              // (Ntot,3)
  const T* img;            // (Ntot,3,2)
  FeatSet<const T> feat;   // n tensors (Ntot,3,C_t): a list of feature tensors is consumed in place (no cat)
  T coord_mult;
  int premultiplied;
  float pix_mult;
  float eps;
  FeatSet<T> out_feat;     // n tensors (B,H,W,C_t)
  int64_t* out_idx;        // (B,H,W)
  T* out_w;                // (B,H,W,3) or nullptr
};

template <typename T>
struct StagedFace {        // SoA in shared memory, one slot per staged face
  uint2 range[256];
  T xy[6][256];
  T z[3][256];
  int id[256];
};

template <typename T>
__global__ void __launch_bounds__(256) raster_fwd_kernel(RasterFwdParams<T> rp, FaceLayout fl, Geometry geo,
                                                         BinLists L) {
  __shared__ StagedFace<T> st;
  __shared__ int scratch[9];

  const int tile = blockIdx.x % geo.tiles_per_view;
  const int b = blockIdx.x / geo.tiles_per_view;
  const int tx = tile % geo.tiles_x, ty = tile / geo.tiles_x;
  const int px = tx * TILE + (threadIdx.x & (TILE - 1));
  const int py = ty * TILE + (threadIdx.x / TILE);
  const bool in_image = px < geo.W && py < geo.H;
  const int64_t vbase = fl.base(b);

  const T x0 = (T)pix_coord(pix_scale(rp.pix_mult, geo.W), geo.W, px);
  const T y0 = (T)pix_coord(pix_scale(rp.pix_mult, geo.H), geo.H, geo.H - 1 - py);

  T best_z = -RN<T>::inf();
  int best_f = -1;
  T bw0 = 0, bw1 = 0, bw2 = 0;
  bool has_nan = false;
  int last_inside = -1;

  const int n_tile = L.tile_count[blockIdx.x];
  const int* tile_list = L.pairs + L.tile_offset[blockIdx.x];
  const int n_large = L.large_count[b];
  const int* large_list = L.large_list + vbase;
  const int rx0 = tx * TILE, ry0 = ty * TILE, rx1 = rx0 + TILE - 1, ry1 = ry0 + TILE - 1;

  const int n_total = n_tile + n_large;
  for (int base = 0; base < n_total; base += 256) {
    // ---- stage up to 256 candidates (tile entries pass unconditionally, large entries are filtered)
    const int k = base + threadIdx.x;
    int f = -1;
    uint2 r = empty_range();
    if (k < n_tile) {
      f = tile_list[k];
      r = L.ranges[vbase + f];
    } else if (k < n_total) {
      f = large_list[k - n_tile];
      r = L.ranges[vbase + f];
      if (!range_hits_rect(r, rx0, ry0, rx1, ry1)) f = -1;
    }
    int staged;
    const int slot = block_compact_256(f >= 0, scratch, staged);   // contains the barriers protecting `st`
    if (f >= 0) {
      T p[6];
      load_face_xy<T>(rp.img, vbase + f, rp.coord_mult, rp.premultiplied, p);
#pragma unroll
      for (int q = 0; q < 6; ++q) st.xy[q][slot] = p[q];
      const T* zz = rp.z + (vbase + f) * 3;
      st.z[0][slot] = zz[0]; st.z[1][slot] = zz[1]; st.z[2][slot] = zz[2];
      st.range[slot] = r;
      st.id[slot] = f;
    }
    __syncthreads();
    // ---- every pixel walks the batch
    if (in_image) {
      for (int s = 0; s < staged; ++s) {
        if (!range_hits_pixel(st.range[s], px, py)) continue;
        T w0, w1, w2;
        if (!bary_weights<T>(st.xy[0][s], st.xy[1][s], st.xy[2][s], st.xy[3][s], st.xy[4][s], st.xy[5][s],
                             x0, y0, rp.eps, w0, w1, w2))
          continue;
        const int fid = st.id[s];
        const T z0 = interp3<T>(st.z[0][s], st.z[1][s], st.z[2][s], w0, w1, w2);
        last_inside = max(last_inside, fid);
        if (z0 != z0) { has_nan = true; continue; }
        if (z0 > best_z || (z0 == best_z && fid < best_f)) {
          best_z = z0; best_f = fid; bw0 = w0; bw1 = w1; bw2 = w2;
        }
      }
    }
    // the next iteration's block_compact_256 starts with a __syncthreads() before `st` is rewritten
  }

  if (!in_image) return;
  if (has_nan) {      // rare: reproduce the in-order scan's outcome (see header)
    best_f = last_inside;
    T p[6];
    load_face_xy<T>(rp.img, vbase + best_f, rp.coord_mult, rp.premultiplied, p);
    bary_weights<T>(p[0], p[1], p[2], p[3], p[4], p[5], x0, y0, rp.eps, bw0, bw1, bw2);
  }
  const int64_t pix = ((int64_t)b * geo.H + py) * geo.W + px;
  rp.out_idx[pix] = (int64_t)best_f;
  if (rp.out_w) {
    T* w = rp.out_w + pix * 3;
    w[0] = bw0; w[1] = bw1; w[2] = bw2;
  }
  for (int t = 0; t < rp.feat.n; ++t) {
    const int C = rp.feat.ch[t];
    T* o = rp.out_feat.ptr[t] + pix * C;
    if (best_f >= 0) {
      const T* ff = rp.feat.ptr[t] + (vbase + best_f) * 3 * C;
      for (int d = 0; d < C; ++d) o[d] = interp3<T>(ff[d], ff[C + d], ff[2 * C + d], bw0, bw1, bw2);
    } else {
      for (int d = 0; d < C; ++d) o[d] = (T)0;
    }
  }
}

}  // namespace dibr
