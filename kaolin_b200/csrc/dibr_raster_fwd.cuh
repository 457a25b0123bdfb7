// dibr_raster_fwd.cuh — K2: per-tile depth-tested face search + feature interpolation.
//
// One CTA = one 16x16 pixel tile of one view.  Two formulations share the epilogue:
//
//  * fp32 — SCATTER (raster_fwd_scatter_kernel), span rasterization.  Each warp takes 32 candidate faces (one per
//    lane, data in registers).  The work items are the (face, pixel row) pairs of the faces' pixel ranges clipped
//    to the tile: flattened over the warp (prefix sum, owner lane by REDUX/ballot/popc, face data by shuffle), every
//    lane derives a CONSERVATIVE column span for its row from the three edge functions (error-bounded margins, see
//    the comment above HIT_CAP and DESIGN.md section 4.5).  The span's pixels are pushed on a per-warp stack in shared
//    memory and popped in full batches of 32 for the reference's exact test (three IEEE divisions), depth, and a
//    CAS-max of (z0, face) into the pixel's 64-bit key.  The cost follows the triangles' area, not their bounding
//    boxes, and the divisions run with all lanes busy.
//  * fp64, and fp32 tiles that meet a NaN depth — GATHER (gather_tile): one thread per pixel walks the
//    staged candidates; reproduces the reference's in-order scan including its NaN behaviour.
//
// Winner rule (order independent, equals the reference's in-order scan rasterization_cuda.cu:88-171):
//   * among faces whose z0 is not NaN: the largest z0 (> -inf), lowest face index on ties;
//   * a NaN never compares: `z0 <= max_z0` (:162) is false both for a NaN z0 (the face is accepted and max_z0 becomes
//     NaN) and for any z0 against a NaN max_z0 (the next inside face is accepted whatever its depth, after which the
//     normal compare resumes).  With n = the largest index among the inside faces with a NaN depth, the winner is
//     therefore the normal rule restricted to the inside faces of index > n (any depth, -inf included, beats "none"),
//     or n itself when there are none — resolve_after_nan() below, a second per-pixel pass taken only by pixels
//     that met a NaN depth.
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
  // dibr_rasterization only (fp32 scatter kernel; nullptr otherwise): the epilogue also writes the soft mask's value of
  // every pixel that the soft-mask kernel may not touch (1 = covered, dibr_soft_mask_cuda.cu:68-70; 0 = uncovered
  // without candidates) and the tile's 256-bit map of uncovered pixels (8 words per tile, bit = t & 31 of word t >> 5,
  // t = row * 16 + column), so that the soft-mask kernel neither re-reads face_idx nor visits finished tiles.
  T* soft_out;             // (B,H,W) or nullptr
  unsigned* tile_unc;      // [B * tiles_per_view * 8] or nullptr
};

template <typename T>
struct StagedFace {        // SoA in shared memory, one slot per staged face
  uint2 range[256];
  T xy[6][256];
  T z[3][256];
  int id[256];
};

// ---------------------------------------------------------------------------------- NaN depths (rare)
// One pixel that met an inside face with a NaN depth, n = the largest index of such a face: the reference's in-order
// scan (:106-170) ends with the best face among the inside faces of index > n — largest z0, lowest index on ties,
// the first of them accepted whatever its depth — or with n itself.  Walks the tile's (unordered) candidate lists.
template <typename T>
__device__ __forceinline__ void resolve_after_nan(const RasterFwdParams<T>& rp, const BinLists& L, const TileLists& tl,
                                               const int64_t vbase, const int px, const int py, const T x0,
                                               const T y0, const int n, int& best_f, T& bw0, T& bw1, T& bw2) {
  T best_z = 0;
  best_f = -1;
  const int n_total = tl.n_fine + tl.n_filtered();
  for (int k = 0; k < n_total; ++k) {
    const int f = k < tl.n_fine ? tl.fine[k] : tl.filtered_at(k - tl.n_fine);
    if (f <= n || !range_hits_pixel(L.ranges[vbase + f], px, py)) continue;
    T p[6], w0, w1, w2;
    load_face_xy<T>(rp.img, vbase + f, rp.coord_mult, rp.premultiplied, p);
    if (!bary_weights<T>(p[0], p[1], p[2], p[3], p[4], p[5], x0, y0, rp.eps, w0, w1, w2)) continue;
    const T* zz = rp.z + (vbase + f) * 3;
    const T z0 = interp3<T>(zz[0], zz[1], zz[2], w0, w1, w2);           // not NaN: f > n
    if (best_f < 0 || z0 > best_z || (z0 == best_z && f < best_f)) { best_z = z0; best_f = f; }
  }
  if (best_f < 0) best_f = n;
  T p[6];
  load_face_xy<T>(rp.img, vbase + best_f, rp.coord_mult, rp.premultiplied, p);
  bary_weights<T>(p[0], p[1], p[2], p[3], p[4], p[5], x0, y0, rp.eps, bw0, bw1, bw2);
}

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
  int nan_idx = -1;                 // largest index among the inside faces with a NaN depth

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
        if (z0 != z0) { nan_idx = max(nan_idx, fid); continue; }
        if (z0 > best_z || (z0 == best_z && fid < best_f)) {
          best_z = z0; best_f = fid; bw0 = w0; bw1 = w1; bw2 = w2;
        }
      }
    }
    // the next iteration's block_compact_256 starts with a __syncthreads() before `st` is rewritten
  }
  if (in_image && nan_idx >= 0)     // rare: reproduce the in-order scan's outcome (see header)
    resolve_after_nan<T>(rp, L, tl, vbase, px, py, x0, y0, nan_idx, best_f, bw0, bw1, bw2);
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

// Feature store of one pixel with a compile-time channel count (the scatter kernel's epilogue dispatches on the
// tensor's C once per tensor, outside its pixel loop: no per-channel loop control or address arithmetic).
template <typename T, int CC>
__device__ __forceinline__ void store_features(const T* __restrict__ feat, T* __restrict__ out, const int64_t vbase,
                                               const int64_t pix, const int best_f, const T w0, const T w1,
                                               const T w2) {
  T* o = out + pix * CC;
  T v[CC];
  if (best_f >= 0) {
    const T* f0 = feat + (vbase + best_f) * 3 * CC;
    interp_block<T, CC>(f0, f0 + CC, f0 + 2 * CC, v, w0, w1, w2);
  } else {
#pragma unroll
    for (int d = 0; d < CC; ++d) v[d] = (T)0;
  }
  // a pixel's channels are contiguous: 2 / 4 fp32 channels go out as one 8 / 16-byte store (STG.E.64 / STG.E.128) when
  // the tensor's base is that aligned (torch allocations are); a warp then writes 2 x 16 pixels = two contiguous
  // 128 / 256-byte row segments with one instruction
  if (sizeof(T) == 4 && CC == 2 && (reinterpret_cast<uintptr_t>(out) & 7u) == 0) {
    *reinterpret_cast<float2*>(o) = make_float2((float)v[0], (float)v[1]);
  } else if (sizeof(T) == 4 && CC == 4 && (reinterpret_cast<uintptr_t>(out) & 15u) == 0) {
    *reinterpret_cast<float4*>(o) = make_float4((float)v[0], (float)v[1], (float)v[CC > 2 ? 2 : 0], (float)v[CC > 3 ? 3 : 0]);
  } else {
#pragma unroll
    for (int d = 0; d < CC; ++d) o[d] = v[d];
  }
}

template <typename T>
__global__ void __launch_bounds__(256) raster_fwd_kernel(RasterFwdParams<T> rp, FaceLayout fl, Geometry geo,
                                                         BinLists L) {
  __shared__ StagedFace<T> st;
  __shared__ int scratch[9];
  int b, tx, ty;
  if (!scheduled_tile(L, geo, b, tx, ty)) return;
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

// ---- span rasterization (fp32).  Work items are (face, pixel row) pairs of the face's pixel range clipped to the
// tile.  For each row a CONSERVATIVE column span is derived from the three edge functions; only the pixels of the
// span ("candidates") get the reference's exact test with its IEEE divisions (resolve_hits), so the cost follows the
// triangle's area instead of its bounding box (sliver meshes: < 1 % of the bbox pixels are inside).
//
// Why no inside pixel can be lost.  With u = 2^-24, the reference's w_i (rasterization_cuda.cu:131-139) differs
// from the exact edge function W_i(x0,y0) by at most 5u S, S = |bex||cey| + |bey||cex| <= 2 Dx Dy, where Dx / Dy
// bound the distance of every vertex from every pixel centre of the clipped range.  W_0 + W_1 + W_2 = N (twice the
// signed area) for every pixel, the reference's norm differs from N by at most 42u Dx Dy and the fp32 estimate nc
// below by at most 40u Dx Dy: if |nc| > 128u Dx Dy (+2^-38 for underflow) the reference's norm has the sign s of nc
// at every pixel of the range (and |norm| <= 2^80 as Dx Dy, |eps| <= 1e18).  A pixel with s W_i <= -(5u S + 2^-39)
// then has s w_i <= -2^-39: a non-zero quotient of the opposite sign, i.e. the reference rejects it.  s W_i is
// linear along a row: t_r + a_i (x0 - xr) with t_r its value at the range's first column, which the row item
// evaluates with the reference's own expression (error <= 5u S), a_i = s dW_i/dx.  Excluding the columns
// d < (t_r + m2) * (-1 / (a_i pitch)) - 1/64 (a_i > 0; mirrored for a_i < 0) with m2 = 32u Dx Dy + 2^-38 leaves a
// margin of 5u S for t_r, 5u S + 2^-39 for the claim and 1/64 pixel for the rounding of the bound (<= 2^-16 px)
// and for the deviation of the fp32 pixel centres from an exact lattice (<= W u <= 2^-9 px for W <= 32768).
// Faces that fail the |nc| test (degenerate or huge) keep their whole range: span = bbox, as before.
constexpr int HIT_CAP = 32 + 32 * TILE;     // < 32 left over + one row iteration's candidates (32 rows x 16 columns)

template <int NW>
struct ScatterShared {
  unsigned long long key[TILE_PIX];
  float px[TILE], py[TILE];       // pixel-centre coordinates of the tile's columns / rows
  int wq[NW][32];                 // per-warp compaction queue for the filtered "large" faces
  unsigned short hit[NW][HIT_CAP];// per-warp stack of candidates: owner lane << 8 | row << 4 | column
  float pitch;                    // x distance of neighbouring pixel centres (2 * multiplier / W)
  int nan_seen;
};

struct FaceRegs {                 // one face per lane (multiplied coordinates), read by shuffle from its owner lane
  float ax, ay, bx, by, cx, cy, az, bz, cz;
  int f;
};

// Exact evaluation of up to 32 candidates (lane < n takes hit[lane]): the reference's barycentric weights with
// their three IEEE divisions, inside test, depth, and the order-independent merge into the pixel's key.  Called
// with all 32 lanes (the face data comes by shuffle from the lane that owns the face).
template <int NW>
__device__ __forceinline__ void resolve_hits(const RasterFwdParams<float>& rp, ScatterShared<NW>& sh,
                                             const FaceRegs& fr, const unsigned short* hit, const int n) {
  const int lane = lane_id();
  const bool act = lane < n;
  const unsigned e = act ? (unsigned)hit[lane] : 0u;
  const int owner = (int)(e >> 8);
  const float ax = __shfl_sync(0xffffffffu, fr.ax, owner), ay = __shfl_sync(0xffffffffu, fr.ay, owner);
  const float bx = __shfl_sync(0xffffffffu, fr.bx, owner), by = __shfl_sync(0xffffffffu, fr.by, owner);
  const float cx = __shfl_sync(0xffffffffu, fr.cx, owner), cy = __shfl_sync(0xffffffffu, fr.cy, owner);
  const float az = __shfl_sync(0xffffffffu, fr.az, owner), bz = __shfl_sync(0xffffffffu, fr.bz, owner);
  const float cz = __shfl_sync(0xffffffffu, fr.cz, owner);
  const int f = __shfl_sync(0xffffffffu, fr.f, owner);
  if (!act) return;
  float w0, w1, w2;
  if (!bary_weights<float>(ax, ay, bx, by, cx, cy, sh.px[e & 15u], sh.py[(e >> 4) & 15u], rp.eps, w0, w1, w2)) return;
  float z0 = interp3<float>(az, bz, cz, w0, w1, w2);
  if (z0 != z0) { sh.nan_seen = 1; return; }                       // tile falls back to the in-order rule
  z0 = __fadd_rn(z0, 0.0f);                                        // -0 -> +0: the reference compares values
  const unsigned long long key = ((unsigned long long)ord_f32(z0) << 32) | (unsigned)(0xffffffffu - (unsigned)f);
  smem_max_u64(&sh.key[e & 255u], key);
}

// One warp, up to 32 faces (lane = face, f < 0 = none; the valid lanes must be the low lanes).
template <int NW>
__device__ __forceinline__ void scatter_faces(const RasterFwdParams<float>& rp, const BinLists& L,
                                              ScatterShared<NW>& sh, const int64_t vbase, const int f,
                                              const int rx0, const int ry0, unsigned short* hit) {
  typedef RN<float> R;
  const int lane = lane_id();
  FaceRegs fr;
  fr.ax = fr.ay = fr.bx = fr.by = fr.cx = fr.cy = fr.az = fr.bz = fr.cz = 0.0f;
  fr.f = f;
  float exa = 0, exb = 0, exc = 0;       // vertex x minus the x of the range's first column (reference rounding)
  float rn0 = 0, rn1 = 0, rn2 = 0;       // -1 / (a_i * pitch); 0 = edge i gives no bound
  float ms = 0;                          // margin m2 carrying the sign s of the face's area
  unsigned rect = 0;
  int h = 0;
  if (f >= 0) {
    const uint2 r = L.ranges[vbase + f];
    const int lx0 = max((int)(r.x & 0xffffu), rx0) - rx0, lx1 = min((int)(r.x >> 16), rx0 + TILE - 1) - rx0;
    const int ly0 = max((int)(r.y & 0xffffu), ry0) - ry0, ly1 = min((int)(r.y >> 16), ry0 + TILE - 1) - ry0;
    const int w = lx1 - lx0 + 1, hh = ly1 - ly0 + 1;
    if (w > 0 && hh > 0) {
      h = hh;
      rect = (unsigned)lx0 | ((unsigned)ly0 << 4) | ((unsigned)(w - 1) << 8);
      float p[6];
      load_face_xy<float>(rp.img, vbase + f, rp.coord_mult, rp.premultiplied, p);
      fr.ax = p[0]; fr.ay = p[1]; fr.bx = p[2]; fr.by = p[3]; fr.cx = p[4]; fr.cy = p[5];
      const float* zz = rp.z + (vbase + f) * 3;
      fr.az = __ldg(zz); fr.bz = __ldg(zz + 1); fr.cz = __ldg(zz + 2);
      const float xl = sh.px[lx0], xh = sh.px[lx1], yl = sh.py[ly0], yh = sh.py[ly1];
      exa = R::sub(p[0], xl); exb = R::sub(p[2], xl); exc = R::sub(p[4], xl);
      const float dx = fmaxf(fmaxf(fmaxf(fabsf(p[0] - xl), fabsf(p[0] - xh)), fmaxf(fabsf(p[2] - xl), fabsf(p[2] - xh))),
                             fmaxf(fabsf(p[4] - xl), fabsf(p[4] - xh)));
      const float dy = fmaxf(fmaxf(fmaxf(fabsf(p[1] - yl), fabsf(p[1] - yh)), fmaxf(fabsf(p[3] - yl), fabsf(p[3] - yh))),
                             fmaxf(fabsf(p[5] - yl), fabsf(p[5] - yh)));
      const float dd = dx * dy;
      const float nc = fmaf(p[2] - p[0], p[5] - p[1], -((p[3] - p[1]) * (p[4] - p[0])));
      if (fabsf(nc) > fmaf(dd, 7.62939453125e-6f /* 2^-17 */, 3.637978807091713e-12f /* 2^-38 */) && dd <= 1e18f &&
          fabsf(rp.eps) <= 1e18f) {                                  // false for NaN / inf operands
        ms = copysignf(fmaf(dd, 1.9073486328125e-6f /* 2^-19 */, 3.637978807091713e-12f), nc);
        const float sp = copysignf(sh.pitch, nc);                    // s * pitch
        const float a0 = (p[3] - p[5]) * sp, a1 = (p[5] - p[1]) * sp, a2 = (p[1] - p[3]) * sp;
        // approximate reciprocals (<= 2 ulp): the bounds they scale are clamped to [-2, 32] columns and carry a
        // 1/64-pixel margin, 2.4e-7 relative is 8e-6 of a column at most (three IEEE divisions per staged face before)
        rn0 = a0 != 0.0f ? __fdividef(-1.0f, a0) : 0.0f;
        rn1 = a1 != 0.0f ? __fdividef(-1.0f, a1) : 0.0f;
        rn2 = a2 != 0.0f ? __fdividef(-1.0f, a2) : 0.0f;
      }
    }
  }
  // exclusive prefix of the row counts
  int incl = h;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const int start = incl - h;
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  const unsigned nonempty = __ballot_sync(0xffffffffu, h > 0);
  const bool holes = (nonempty & (nonempty + 1u)) != 0u;          // non-empty lanes are not a prefix (never by
                                                                  // construction; handled for robustness)
  const unsigned le_mask = 0xffffffffu >> (31 - lane);
  int nhit = 0;                                                     // warp-uniform stack height
  for (int c0 = 0; c0 < total; c0 += 32) {
    const unsigned rel = (unsigned)(start - c0);
    const unsigned bit = (h > 0 && rel < 32u) ? (1u << rel) : 0u;
    const unsigned starts = __reduce_or_sync(0xffffffffu, bit);
    const int before = __popc(__ballot_sync(0xffffffffu, h > 0 && start < c0));
    int owner = before + __popc(starts & le_mask) - 1;             // rank among the non-empty faces
    if (holes) owner = __fns(nonempty, 0, owner + 1);
    const int p = c0 + lane;
    const bool active = p < total;
    owner = active ? owner : 0;
    const int ostart = __shfl_sync(0xffffffffu, start, owner);
    const unsigned orect = __shfl_sync(0xffffffffu, rect, owner);
    const float oay = __shfl_sync(0xffffffffu, fr.ay, owner), oby = __shfl_sync(0xffffffffu, fr.by, owner);
    const float ocy = __shfl_sync(0xffffffffu, fr.cy, owner);
    const float oxa = __shfl_sync(0xffffffffu, exa, owner), oxb = __shfl_sync(0xffffffffu, exb, owner);
    const float oxc = __shfl_sync(0xffffffffu, exc, owner);
    const float on0 = __shfl_sync(0xffffffffu, rn0, owner), on1 = __shfl_sync(0xffffffffu, rn1, owner);
    const float on2 = __shfl_sync(0xffffffffu, rn2, owner);
    const float oms = __shfl_sync(0xffffffffu, ms, owner);
    int len = 0;
    unsigned ebase = 0;
    if (active) {
      const unsigned ly = (((orect >> 4) & 15u) + (unsigned)(p - ostart)) & 15u;
      const int wm1 = (int)((orect >> 8) & 15u);
      const float y0 = sh.py[ly];
      const float aey = R::sub(oay, y0), bey = R::sub(oby, y0), cey = R::sub(ocy, y0);
      const float w0 = R::fma(oxb, cey, -R::mul(bey, oxc));          // the reference's w_i at the range's first column
      const float w1 = R::fma(oxc, aey, -R::mul(cey, oxa));
      const float w2 = R::fma(oxa, bey, -R::mul(aey, oxb));
      const unsigned sbit = __float_as_uint(oms) & 0x80000000u;
      const float m2 = fabsf(oms);
      const float l0 = (__uint_as_float(__float_as_uint(w0) ^ sbit) + m2) * on0;
      const float l1 = (__uint_as_float(__float_as_uint(w1) ^ sbit) + m2) * on1;
      const float l2 = (__uint_as_float(__float_as_uint(w2) ^ sbit) + m2) * on2;
      float lof = 0.0f, hif = (float)wm1;                            // NaN bounds fail every compare: no constraint
      if (on0 < 0.0f && l0 > lof) lof = l0;
      if (on0 > 0.0f && l0 < hif) hif = l0;
      if (on1 < 0.0f && l1 > lof) lof = l1;
      if (on1 > 0.0f && l1 < hif) hif = l1;
      if (on2 < 0.0f && l2 > lof) lof = l2;
      if (on2 > 0.0f && l2 < hif) hif = l2;
      lof = fminf(lof, 32.0f); hif = fmaxf(hif, -2.0f);
      const int dmin = max((int)ceilf(lof - 0.015625f), 0);
      const int dmax = min((int)floorf(hif + 0.015625f), wm1);
      len = max(dmax - dmin + 1, 0);
      ebase = ((unsigned)owner << 8) | (ly << 4) | ((orect & 15u) + (unsigned)dmin);
    }
    // push the candidates: exclusive prefix of the span lengths, then every lane writes its own
    int inc2 = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc2, o);
      if (lane >= o) inc2 += t;
    }
    const int tot2 = __shfl_sync(0xffffffffu, inc2, 31);
    unsigned short* dst = hit + nhit + (inc2 - len);
    for (int j = 0; j < len; ++j) dst[j] = (unsigned short)(ebase + (unsigned)j);
    nhit += tot2;
    __syncwarp();
    while (nhit >= 32) {                                             // full batches off the top of the stack
      nhit -= 32;
      resolve_hits<NW>(rp, sh, fr, hit + nhit, 32);
    }
    __syncwarp();
  }
  if (nhit > 0) {                                                    // the faces change with the next call
    resolve_hits<NW>(rp, sh, fr, hit, nhit);
    __syncwarp();
  }
}

// Rare fallback of the scatter kernel for tiles that met a NaN depth: the reference's in-order scan restated per
// pixel over the tile's (unordered) candidate lists with the order-independent rule of the file header.
__device__ __forceinline__ void scan_pixel_in_order_rule(const RasterFwdParams<float>& rp, const BinLists& L,
                                                         const TileLists& tl, const int64_t vbase, const int px,
                                                         const int py, const float x0, const float y0, int& best_f,
                                                         float& bw0, float& bw1, float& bw2) {
  float best_z = -RN<float>::inf();
  int nan_idx = -1;
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
    if (z0 != z0) { nan_idx = max(nan_idx, f); continue; }
    if (z0 > best_z || (z0 == best_z && f < best_f)) { best_z = z0; best_f = f; bw0 = w0; bw1 = w1; bw2 = w2; }
  }
  if (nan_idx >= 0) resolve_after_nan<float>(rp, L, tl, vbase, px, py, x0, y0, nan_idx, best_f, bw0, bw1, bw2);
}

// NT threads per CTA (one CTA per 16x16 tile whatever NT).
#ifndef DIBR_RASTER_TPS
#define DIBR_RASTER_TPS 1024       // resident threads per SM the register allocation aims at
#endif
template <int NT>
__global__ void __launch_bounds__(NT, DIBR_RASTER_TPS / NT)
raster_fwd_scatter_kernel(RasterFwdParams<float> rp, FaceLayout fl, Geometry geo, BinLists L) {
  constexpr int NW = NT / 32;
  __shared__ ScatterShared<NW> sh;
  int b, tx, ty;
  if (!scheduled_tile(L, geo, b, tx, ty)) return;
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
  if (threadIdx.x == 0) {
    sh.nan_seen = 0;
    sh.pitch = __fmul_rn(2.0f, pix_scale(rp.pix_mult, geo.W));
  }
  __syncthreads();

  const TileLists tl = tile_lists(L, geo, fl, b, tx, ty);
  const int n_tile = tl.n_fine;
  const int* tile_list = tl.fine;
  // the tile's faces are dealt to the warps in equal contiguous slices (a short list still keeps every warp
  // busy: the parallelism is in the pairs, not in the faces)
  const int per_warp = (n_tile + NW - 1) / NW;
  const int slice_end = min(n_tile, (warp + 1) * per_warp);
  unsigned short* hit = sh.hit[warp];
  for (int base = warp * per_warp; base < slice_end; base += 32) {   // warp-uniform trip count
    const int k = base + lane;
    scatter_faces<NW>(rp, L, sh, vbase, (k < slice_end) ? tile_list[k] : -1, rx0, ry0, hit);
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
    scatter_faces<NW>(rp, L, sh, vbase, f, rx0, ry0, hit);
  }
  __syncthreads();

  const bool nan_seen = sh.nan_seen != 0;                             // CTA-uniform, rare
  unsigned* const unc = rp.tile_unc
      ? rp.tile_unc + ((size_t)b * geo.tiles_per_view + (size_t)(ty * geo.tiles_x + tx)) * (TILE_PIX / 32) : nullptr;
  for (int t = threadIdx.x; t < TILE_PIX; t += NT) {                   // uniform trip count (TILE_PIX % NT == 0)
    const int px = rx0 + (t & (TILE - 1)), py = ry0 + (t / TILE);
    const bool in_image = px < geo.W && py < geo.H;
    int best_f = -1;
    if (in_image) {
      const float x0 = sh.px[t & (TILE - 1)], y0 = sh.py[t / TILE];
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
      const int64_t pix = ((int64_t)b * geo.H + py) * geo.W + px;
      rp.out_idx[pix] = (int64_t)best_f;
      if (rp.soft_out) rp.soft_out[pix] = best_f >= 0 ? 1.0f : 0.0f;    // final unless the soft-mask kernel finds hits
      if (rp.out_w) {
        float* w = rp.out_w + pix * 3;
        w[0] = bw0; w[1] = bw1; w[2] = bw2;
      }
      // rasterization_cuda.cu:172-189 per feature tensor; small channel counts take fully unrolled loads / stores
      for (int ft = 0; ft < rp.feat.n; ++ft) {
        const int C = rp.feat.ch[ft];
        const float* feat = rp.feat.ptr[ft];
        float* out = rp.out_feat.ptr[ft];
        switch (C) {
          case 1: store_features<float, 1>(feat, out, vbase, pix, best_f, bw0, bw1, bw2); break;
          case 2: store_features<float, 2>(feat, out, vbase, pix, best_f, bw0, bw1, bw2); break;
          case 3: store_features<float, 3>(feat, out, vbase, pix, best_f, bw0, bw1, bw2); break;
          case 4: store_features<float, 4>(feat, out, vbase, pix, best_f, bw0, bw1, bw2); break;
          default: {
            float* o = out + pix * C;
            if (best_f >= 0) {
              const float* f0 = feat + (vbase + best_f) * 3 * C;
              const float* f1 = f0 + C;
              const float* f2 = f1 + C;
              int d = 0;
              for (; d + 4 <= C; d += 4) interp_block<float, 4>(f0 + d, f1 + d, f2 + d, o + d, bw0, bw1, bw2);
              if (d + 2 <= C) { interp_block<float, 2>(f0 + d, f1 + d, f2 + d, o + d, bw0, bw1, bw2); d += 2; }
              if (d < C) interp_block<float, 1>(f0 + d, f1 + d, f2 + d, o + d, bw0, bw1, bw2);
            } else {
              for (int d = 0; d < C; ++d) o[d] = 0.0f;
            }
          }
        }
      }
    }
    if (unc) {                                                         // (CTA-uniform)
      const unsigned um = __ballot_sync(0xffffffffu, in_image && best_f < 0);
      if (lane == 0) unc[t >> 5] = um;
    }
  }
}

inline int raster_threads_default() {
  static const int nt = [] {
    const char* on = getenv("DIBR_B200_TUNING");       // the one documented switch (include/dibr_b200.h)
    const char* e = (on && on[0] == '1' && on[1] == 0) ? getenv("DIBR_RASTER_NT") : nullptr;
    const int v = e ? atoi(e) : 0;
    return (v == 64 || v == 128 || v == 256) ? v : 64;
  }();
  return nt;
}

// dispatch: fp32 -> scatter, fp64 -> gather
inline void launch_raster_fwd(const RasterFwdParams<float>& rp, const FaceLayout& fl, const Geometry& geo,
                              const BinLists& L, cudaStream_t s) {
  const int nt = raster_threads_default();          // DIBR_RASTER_NT=64|128|256 overrides for experiments
  if (nt == 64) raster_fwd_scatter_kernel<64><<<scheduled_grid(geo), 64, 0, s>>>(rp, fl, geo, L);
  else if (nt == 128) raster_fwd_scatter_kernel<128><<<scheduled_grid(geo), 128, 0, s>>>(rp, fl, geo, L);
  else raster_fwd_scatter_kernel<256><<<scheduled_grid(geo), 256, 0, s>>>(rp, fl, geo, L);
}
inline void launch_raster_fwd(const RasterFwdParams<double>& rp, const FaceLayout& fl, const Geometry& geo,
                              const BinLists& L, cudaStream_t s) {
  raster_fwd_kernel<double><<<scheduled_grid(geo), 256, 0, s>>>(rp, fl, geo, L);
}

}  // namespace dibr
