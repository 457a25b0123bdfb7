// dibr_raster_fwd.cuh — K2: per-tile depth-tested face search + feature interpolation.
//
// One CTA = one 16x16 pixel tile of one view.  Two formulations share the epilogue:
//
//  * fp32 — SCATTER (raster_fwd_scatter_kernel).  The work items are (face, pixel) pairs with the pixel
//    inside the face's pixel range clipped to the tile.  Each warp takes 32 candidate faces (one per lane,
//    data in registers), prefix-sums their pair counts and then walks the flattened pair list 32 pairs at
//    a time: owner lane by REDUX/ballot/popc, face data by shuffle.  Every lane evaluates one pair with the
//    reference's barycentric arithmetic (cheap exact sign rejection first, the three IEEE divisions only
//    for pixels that can be inside) and merges (z0, face) into the pixel's 64-bit key in shared memory
//    with a CAS-max.  Lanes stay full whatever the triangle size; no per-pixel walk over faces the pixel
//    is not in.
//  * fp64, and fp32 tiles that meet a NaN depth — GATHER (gather_tile): one thread per pixel walks the
//    staged candidates; reproduces the reference's in-order scan including its NaN behaviour.
//
// Winner rule (order independent, equals the reference's in-order scan rasterization_cuda.cu:88-171):
//   * among faces whose z0 is not NaN: the largest z0 (> -inf), lowest face index on ties;
//   * if any inside face has a NaN z0 the reference's `z0 <= max_z0` is false from then on, so the
//     LAST inside face in index order wins — tracked as (has_nan, max inside index) by the gather path.
#pragma once
#include "dibr_binning.cuh"
#include <cstdlib>

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

// ---------------------------------------------------------------------------------- gather (per pixel)
template <typename T>
__device__ __forceinline__ void gather_tile(const RasterFwdParams<T>& rp, const FaceLayout& fl, const Geometry& geo,
                                            const BinLists& L, StagedFace<T>& st, int* scratch, const int b,
                                            const int tx, const int ty, const int px, const int py,
                                            const bool in_image, const T x0, const T y0, int& best_f, T& bw0,
                                            T& bw1, T& bw2) {
  const int64_t vbase = fl.base(b);
  T best_z = -RN<T>::inf();
  best_f = -1;
  bw0 = 0; bw1 = 0; bw2 = 0;
  bool has_nan = false;
  int last_inside = -1;

  const TileLists tl = tile_lists(L, geo, fl, b, tx, ty);
  const int n_tile = tl.n_fine;
  const int rx0 = tx * TILE, ry0 = ty * TILE, rx1 = rx0 + TILE - 1, ry1 = ry0 + TILE - 1;

  const int n_total = n_tile + tl.n_filtered();
  for (int base = 0; base < n_total; base += 256) {
    // ---- stage up to 256 candidates (tile entries pass unconditionally, coarse / large entries are filtered)
    const int k = base + threadIdx.x;
    int f = -1;
    uint2 r = empty_range();
    if (k < n_tile) {
      f = tl.fine[k];
      r = L.ranges[vbase + f];
    } else if (k < n_total) {
      f = tl.filtered_at(k - n_tile);
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
  if (in_image && has_nan) {      // rare: reproduce the in-order scan's outcome (see header)
    best_f = last_inside;
    T p[6];
    load_face_xy<T>(rp.img, vbase + best_f, rp.coord_mult, rp.premultiplied, p);
    bary_weights<T>(p[0], p[1], p[2], p[3], p[4], p[5], x0, y0, rp.eps, bw0, bw1, bw2);
  }
}

// ---------------------------------------------------------------------------------- epilogue (per pixel)
// rasterization_cuda.cu:172-189 + the packed->original index mapping of rasterization.py:357-363.
// Channels are handled in compile-time blocks of 4/2/1 so that the 3*CC loads and CC stores of a block use
// immediate offsets from three vertex pointers (no per-channel address arithmetic).
template <typename T, int CC>
__device__ __forceinline__ void interp_block(const T* __restrict__ f0, const T* __restrict__ f1,
                                             const T* __restrict__ f2, T* __restrict__ o, const T w0, const T w1,
                                             const T w2) {
  T a[CC], b[CC], c[CC];
#pragma unroll
  for (int d = 0; d < CC; ++d) { a[d] = __ldg(f0 + d); b[d] = __ldg(f1 + d); c[d] = __ldg(f2 + d); }
#pragma unroll
  for (int d = 0; d < CC; ++d) o[d] = interp3<T>(a[d], b[d], c[d], w0, w1, w2);
}
template <typename T>
__device__ __forceinline__ void raster_epilogue(const RasterFwdParams<T>& rp, const int64_t vbase, const int64_t pix,
                                                const int best_f, const T bw0, const T bw1, const T bw2) {
  rp.out_idx[pix] = (int64_t)best_f;
  if (rp.out_w) {
    T* w = rp.out_w + pix * 3;
    w[0] = bw0; w[1] = bw1; w[2] = bw2;
  }
  for (int t = 0; t < rp.feat.n; ++t) {
    const int C = rp.feat.ch[t];
    T* o = rp.out_feat.ptr[t] + pix * C;
    if (best_f >= 0) {
      const T* f0 = rp.feat.ptr[t] + (vbase + best_f) * 3 * C;
      const T* f1 = f0 + C;
      const T* f2 = f1 + C;
      int d = 0;
      for (; d + 4 <= C; d += 4) interp_block<T, 4>(f0 + d, f1 + d, f2 + d, o + d, bw0, bw1, bw2);
      if (d + 2 <= C) { interp_block<T, 2>(f0 + d, f1 + d, f2 + d, o + d, bw0, bw1, bw2); d += 2; }
      if (d < C) interp_block<T, 1>(f0 + d, f1 + d, f2 + d, o + d, bw0, bw1, bw2);
    } else {
      for (int d = 0; d < C; ++d) o[d] = (T)0;
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) raster_fwd_kernel(RasterFwdParams<T> rp, FaceLayout fl, Geometry geo,
                                                         BinLists L) {
  __shared__ StagedFace<T> st;
  __shared__ int scratch[9];
  int b, tile;
  scheduled_tile(L, geo, blockIdx.x, b, tile);
  const int tx = tile % geo.tiles_x, ty = tile / geo.tiles_x;
  const int px = tx * TILE + (threadIdx.x & (TILE - 1));
  const int py = ty * TILE + (threadIdx.x / TILE);
  const bool in_image = px < geo.W && py < geo.H;
  const T x0 = (T)pix_coord(pix_scale(rp.pix_mult, geo.W), geo.W, px);
  const T y0 = (T)pix_coord(pix_scale(rp.pix_mult, geo.H), geo.H, geo.H - 1 - py);
  int best_f;
  T bw0, bw1, bw2;
  gather_tile<T>(rp, fl, geo, L, st, scratch, b, tx, ty, px, py, in_image, x0, y0, best_f, bw0, bw1, bw2);
  if (!in_image) return;
  raster_epilogue<T>(rp, fl.base(b), ((int64_t)b * geo.H + py) * geo.W + px, best_f, bw0, bw1, bw2);
}

// ---------------------------------------------------------------------------------- scatter (fp32)
// 64-bit per-pixel key: high word = depth mapped to an order-preserving unsigned, low word = ~face index,
// so that an unsigned max keeps the largest z0 and, on exact ties, the lowest face index.
__device__ __forceinline__ unsigned ord_f32(float f) {
  const unsigned u = __float_as_uint(f);
  return u ^ ((u & 0x80000000u) ? 0xffffffffu : 0x80000000u);
}
constexpr unsigned ORD_NEG_INF = 0x007fffffu;                         // ord_f32(-inf): z0 must be > -inf to win
constexpr unsigned long long KEY_NONE = ((unsigned long long)ORD_NEG_INF << 32) | 0xffffffffull;

__device__ __forceinline__ void smem_max_u64(unsigned long long* addr, const unsigned long long v) {
  unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(addr);
  while (v > cur) {
    const unsigned long long prev = atomicCAS(addr, cur, v);
    if (prev == cur) break;
    cur = prev;
  }
}

// floor(q / w) for 0 <= q < 256, 1 <= w <= 16 is (q * RECIP16[w]) >> 16 with RECIP16[w] = ceil(65536 / w).
__constant__ unsigned RECIP16[17] = {0, 65536, 32768, 21846, 16384, 13108, 10923, 9363, 8192,
                                     7282, 6554, 5958, 5462, 5042, 4682, 4370, 4096};

constexpr int HIT_BUF = 64;        // per-warp deferred-hit stack: < 32 left over + <= 32 appended per iteration

struct ScatterShared {
  unsigned long long key[TILE_PIX];
  float px[TILE], py[TILE];       // pixel-centre coordinates of the tile's columns / rows
  int wq[8][32];                  // per-warp compaction queue for the filtered "large" faces
  int2 hit[8][HIT_BUF];           // per-warp stack of (face, tile-local pixel) pairs awaiting the exact test
  int nan_seen;
};

// Exact evaluation of up to 32 deferred pairs (lane < n takes hit[lane]): the reference's barycentric weights with
// their three IEEE divisions, inside test, depth, and the order-independent merge into the pixel's key.  All lanes
// are busy here, whereas evaluated in the pair loop the divisions ran for the few inside lanes of each iteration.
__device__ __forceinline__ void resolve_hits(const RasterFwdParams<float>& rp, ScatterShared& sh,
                                             const int64_t vbase, const int2* hit, const int n) {
  const int lane = lane_id();
  if (lane >= n) return;
  const int2 h = hit[lane];
  float p[6], w0, w1, w2;
  load_face_xy<float>(rp.img, vbase + h.x, rp.coord_mult, rp.premultiplied, p);
  const float* zz = rp.z + (vbase + h.x) * 3;
  const float az = __ldg(zz), bz = __ldg(zz + 1), cz = __ldg(zz + 2);
  if (!bary_weights<float>(p[0], p[1], p[2], p[3], p[4], p[5], sh.px[h.y & (TILE - 1)], sh.py[h.y >> 4], rp.eps,
                           w0, w1, w2))
    return;
  float z0 = interp3<float>(az, bz, cz, w0, w1, w2);
  if (z0 != z0) { sh.nan_seen = 1; return; }                       // tile falls back to the in-order rule
  z0 = __fadd_rn(z0, 0.0f);                                        // -0 -> +0: the reference compares values
  const unsigned long long key = ((unsigned long long)ord_f32(z0) << 32) | (unsigned)(0xffffffffu - (unsigned)h.x);
  smem_max_u64(&sh.key[h.y], key);
}

// One warp, up to 32 faces (lane = face, f < 0 = none; the valid lanes must be the low lanes).  `nhit` (warp-
// uniform) = entries on the warp's hit stack; the caller resolves what is left after its last call.
__device__ __forceinline__ void scatter_faces(const RasterFwdParams<float>& rp, const BinLists& L,
                                              ScatterShared& sh, const int64_t vbase, const int f,
                                              const int rx0, const int ry0, int2* hit, int& nhit) {
  const int lane = lane_id();
  float ax = 0, ay = 0, bx = 0, by = 0, cx = 0, cy = 0;
  unsigned rect = 0;
  int n = 0;
  if (f >= 0) {
    const uint2 r = L.ranges[vbase + f];
    const int lx0 = max((int)(r.x & 0xffffu), rx0) - rx0, lx1 = min((int)(r.x >> 16), rx0 + TILE - 1) - rx0;
    const int ly0 = max((int)(r.y & 0xffffu), ry0) - ry0, ly1 = min((int)(r.y >> 16), ry0 + TILE - 1) - ry0;
    const int w = lx1 - lx0 + 1, h = ly1 - ly0 + 1;
    if (w > 0 && h > 0) {
      n = w * h;
      rect = (unsigned)lx0 | ((unsigned)ly0 << 4) | ((unsigned)(w - 1) << 8) | (RECIP16[w] << 12);
      float p[6];
      load_face_xy<float>(rp.img, vbase + f, rp.coord_mult, rp.premultiplied, p);
      ax = p[0]; ay = p[1]; bx = p[2]; by = p[3]; cx = p[4]; cy = p[5];
    }
  }
  // exclusive prefix of the pair counts
  int incl = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const int start = incl - n;
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  const unsigned nonempty = __ballot_sync(0xffffffffu, n > 0);
  const bool holes = (nonempty & (nonempty + 1u)) != 0u;          // non-empty lanes are not a prefix (never by
                                                                  // construction; handled for robustness)
  const unsigned le_mask = 0xffffffffu >> (31 - lane);
  const unsigned lt_mask = le_mask >> 1;
  for (int c0 = 0; c0 < total; c0 += 32) {
    const unsigned rel = (unsigned)(start - c0);
    const unsigned bit = (n > 0 && rel < 32u) ? (1u << rel) : 0u;
    const unsigned starts = __reduce_or_sync(0xffffffffu, bit);
    const int before = __popc(__ballot_sync(0xffffffffu, n > 0 && start < c0));
    int owner = before + __popc(starts & le_mask) - 1;             // rank among the non-empty faces
    if (holes) owner = __fns(nonempty, 0, owner + 1);
    const int p = c0 + lane;
    const bool active = p < total;
    owner = active ? owner : 0;
    const int ostart = __shfl_sync(0xffffffffu, start, owner);
    const unsigned orect = __shfl_sync(0xffffffffu, rect, owner);
    const float oax = __shfl_sync(0xffffffffu, ax, owner), oay = __shfl_sync(0xffffffffu, ay, owner);
    const float obx = __shfl_sync(0xffffffffu, bx, owner), oby = __shfl_sync(0xffffffffu, by, owner);
    const float ocx = __shfl_sync(0xffffffffu, cx, owner), ocy = __shfl_sync(0xffffffffu, cy, owner);
    const int of = __shfl_sync(0xffffffffu, f, owner);
    const unsigned q = (unsigned)(p - ostart);
    const unsigned wq = ((orect >> 8) & 15u) + 1u;
    const unsigned row = (q * (orect >> 12)) >> 16;
    const unsigned col = q - row * wq;
    const unsigned lx = ((orect & 15u) + col) & 15u, ly = (((orect >> 4) & 15u) + row) & 15u;   // (& 15: inactive lanes)
    // division-free: pairs that are not certainly outside go on the warp's stack and get the exact test in full
    // batches of 32 (resolve_hits)
    const bool maybe = active && !bary_certain_reject(oax, oay, obx, oby, ocx, ocy, sh.px[lx], sh.py[ly], rp.eps);
    const unsigned hm = __ballot_sync(0xffffffffu, maybe);
    if (hm != 0u) {                                                  // warp-uniform
      if (maybe) hit[nhit + __popc(hm & lt_mask)] = make_int2(of, (int)(ly * TILE + lx));
      nhit += __popc(hm);
      __syncwarp();
      if (nhit >= 32) {
        nhit -= 32;
        resolve_hits(rp, sh, vbase, hit + nhit, 32);                 // the top 32 entries
        __syncwarp();
      }
    }
  }
}

// Rare fallback of the scatter kernel for tiles that met a NaN depth: the reference's in-order scan restated per
// pixel over the tile's (unordered) candidate lists with the order-independent rule of the file header.
__device__ __forceinline__ void scan_pixel_in_order_rule(const RasterFwdParams<float>& rp, const BinLists& L,
                                                         const TileLists& tl, const int64_t vbase, const int px,
                                                         const int py, const float x0, const float y0, int& best_f,
                                                         float& bw0, float& bw1, float& bw2) {
  float best_z = -RN<float>::inf();
  bool has_nan = false;
  int last_inside = -1;
  best_f = -1; bw0 = 0; bw1 = 0; bw2 = 0;
  const int n_total = tl.n_fine + tl.n_filtered();
  for (int k = 0; k < n_total; ++k) {
    const int f = k < tl.n_fine ? tl.fine[k] : tl.filtered_at(k - tl.n_fine);
    if (!range_hits_pixel(L.ranges[vbase + f], px, py)) continue;
    float p[6], w0, w1, w2;
    load_face_xy<float>(rp.img, vbase + f, rp.coord_mult, rp.premultiplied, p);
    if (!bary_weights<float>(p[0], p[1], p[2], p[3], p[4], p[5], x0, y0, rp.eps, w0, w1, w2)) continue;
    const float* zz = rp.z + (vbase + f) * 3;
    const float z0 = interp3<float>(zz[0], zz[1], zz[2], w0, w1, w2);
    last_inside = max(last_inside, f);
    if (z0 != z0) { has_nan = true; continue; }
    if (z0 > best_z || (z0 == best_z && f < best_f)) { best_z = z0; best_f = f; bw0 = w0; bw1 = w1; bw2 = w2; }
  }
  if (has_nan) {
    best_f = last_inside;
    float p[6];
    load_face_xy<float>(rp.img, vbase + best_f, rp.coord_mult, rp.premultiplied, p);
    bary_weights<float>(p[0], p[1], p[2], p[3], p[4], p[5], x0, y0, rp.eps, bw0, bw1, bw2);
  }
}

// NT threads per CTA (one CTA per 16x16 tile whatever NT).
#ifndef DIBR_RASTER_TPS
#define DIBR_RASTER_TPS 1024       // resident threads per SM the register allocation aims at
#endif
template <int NT>
__global__ void __launch_bounds__(NT, DIBR_RASTER_TPS / NT)
raster_fwd_scatter_kernel(RasterFwdParams<float> rp, FaceLayout fl, Geometry geo, BinLists L) {
  constexpr int NW = NT / 32;
  __shared__ ScatterShared sh;
  int b, tile;
  scheduled_tile(L, geo, blockIdx.x, b, tile);
  const int tx = tile % geo.tiles_x, ty = tile / geo.tiles_x;
  const int rx0 = tx * TILE, ry0 = ty * TILE;
  const int64_t vbase = fl.base(b);
  const int lane = lane_id(), warp = threadIdx.x >> 5;

  for (int t = threadIdx.x; t < TILE_PIX; t += NT) sh.key[t] = KEY_NONE;
  if (threadIdx.x < TILE) {
    sh.px[threadIdx.x] = pix_coord(pix_scale(rp.pix_mult, geo.W), geo.W, rx0 + threadIdx.x);
  } else if (threadIdx.x < 2 * TILE) {
    const int j = threadIdx.x - TILE;
    sh.py[j] = pix_coord(pix_scale(rp.pix_mult, geo.H), geo.H, geo.H - 1 - (ry0 + j));
  }
  if (threadIdx.x == 0) sh.nan_seen = 0;
  __syncthreads();

  const TileLists tl = tile_lists(L, geo, fl, b, tx, ty);
  const int n_tile = tl.n_fine;
  const int* tile_list = tl.fine;
  // the tile's faces are dealt to the warps in equal contiguous slices (a short list still keeps every warp
  // busy: the parallelism is in the pairs, not in the faces)
  const int per_warp = (n_tile + NW - 1) / NW;
  const int slice_end = min(n_tile, (warp + 1) * per_warp);
  int2* hit = sh.hit[warp];
  int nhit = 0;
  for (int base = warp * per_warp; base < slice_end; base += 32) {   // warp-uniform trip count
    const int k = base + lane;
    scatter_faces(rp, L, sh, vbase, (k < slice_end) ? tile_list[k] : -1, rx0, ry0, hit, nhit);
  }
  // faces binned at the coarse level / in the view's large list: filter against the tile, compact, scatter
  const int n_filt = tl.n_filtered();
  for (int base = warp * 32; base < n_filt; base += NT) {
    const int k = base + lane;
    int f = -1;
    if (k < n_filt) {
      f = tl.filtered_at(k);
      if (!range_hits_rect(L.ranges[vbase + f], rx0, ry0, rx0 + TILE - 1, ry0 + TILE - 1)) f = -1;
    }
    const unsigned m = __ballot_sync(0xffffffffu, f >= 0);
    if (m == 0u) continue;
    if (f >= 0) sh.wq[warp][__popc(m & ((1u << lane) - 1u))] = f;     // compact the hits to the low lanes
    __syncwarp();
    f = (lane < __popc(m)) ? sh.wq[warp][lane] : -1;
    __syncwarp();
    scatter_faces(rp, L, sh, vbase, f, rx0, ry0, hit, nhit);
  }
  resolve_hits(rp, sh, vbase, hit, nhit);                             // what is left on the warp's stack (< 32)
  __syncthreads();

  const bool nan_seen = sh.nan_seen != 0;                             // CTA-uniform, rare
  for (int t = threadIdx.x; t < TILE_PIX; t += NT) {
    const int px = rx0 + (t & (TILE - 1)), py = ry0 + (t / TILE);
    if (px >= geo.W || py >= geo.H) continue;
    const float x0 = sh.px[t & (TILE - 1)], y0 = sh.py[t / TILE];
    int best_f = -1;
    float bw0 = 0, bw1 = 0, bw2 = 0;
    if (nan_seen) {
      scan_pixel_in_order_rule(rp, L, tl, vbase, px, py, x0, y0, best_f, bw0, bw1, bw2);
    } else {
      const unsigned long long key = sh.key[t];
      if ((unsigned)(key >> 32) != ORD_NEG_INF) {
        best_f = (int)(0xffffffffu - (unsigned)key);
        float p[6];
        load_face_xy<float>(rp.img, vbase + best_f, rp.coord_mult, rp.premultiplied, p);
        bary_weights<float>(p[0], p[1], p[2], p[3], p[4], p[5], x0, y0, rp.eps, bw0, bw1, bw2);
      }
    }
    raster_epilogue<float>(rp, vbase, ((int64_t)b * geo.H + py) * geo.W + px, best_f, bw0, bw1, bw2);
  }
}

inline int raster_threads_default() {
  static const int nt = [] {
    const char* e = getenv("DIBR_RASTER_NT");
    const int v = e ? atoi(e) : 0;
    return (v == 64 || v == 128 || v == 256) ? v : 128;
  }();
  return nt;
}

// dispatch: fp32 -> scatter, fp64 -> gather
inline void launch_raster_fwd(const RasterFwdParams<float>& rp, const FaceLayout& fl, const Geometry& geo,
                              const BinLists& L, unsigned grid, cudaStream_t s) {
  const int nt = raster_threads_default();          // DIBR_RASTER_NT=64|128|256 overrides for experiments
  if (nt == 64) raster_fwd_scatter_kernel<64><<<grid, 64, 0, s>>>(rp, fl, geo, L);
  else if (nt == 128) raster_fwd_scatter_kernel<128><<<grid, 128, 0, s>>>(rp, fl, geo, L);
  else raster_fwd_scatter_kernel<256><<<grid, 256, 0, s>>>(rp, fl, geo, L);
}
inline void launch_raster_fwd(const RasterFwdParams<double>& rp, const FaceLayout& fl, const Geometry& geo,
                              const BinLists& L, unsigned grid, cudaStream_t s) {
  raster_fwd_kernel<double><<<grid, 256, 0, s>>>(rp, fl, geo, L);
}

}  // namespace dibr
