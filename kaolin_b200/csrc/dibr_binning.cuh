// dibr_binning.cuh — per-face setup (K0) and tile binning (K1) shared by the rasterizer and the
// soft-mask kernels.
//
// K0/K1a  setup_count : one thread per face.  Replaces the torch glue of the reference
//         (rasterization.py:309-344 valid-face packing + x multiplier + bbox; dibr.py:32-39 enlarged bbox)
//         with one coalesced pass: multiplied coordinates -> bbox -> EXACT inclusive pixel range of the
//         reference's half-open float test (rasterization_cuda.cu:115) -> 8-byte range record, and
//         counts the face into every 16x16 tile it can touch.
// K1b     scan        : list offsets.  Every 1024 counters are scanned by their own CTA, whose segment base comes from
//         one atomicAdd on the allocation cursor: a tile's list only has to be contiguous in itself, not with its
//         neighbours', so no cross-CTA prefix (no look-back) is needed and a view's counters are spread over
//         several SMs instead of one.  Also counts, per view, the tiles in every log2 weight bucket.
// K1c     fill        : scatter face ids into the per-tile lists, and write each view's tile launch order
//         (heaviest bucket first) from the bucket counts.
// One call can build TWO sets of lists from a single read of the coordinates (`DUAL`): the rasterizer's (tight
// bboxes, valid faces only) and the soft mask's (bboxes enlarged by boxlen * multiplier, all faces) — the
// dibr_rasterization entry point; 3 launches + 1 memset instead of 6 + 2.
//
// Two levels.  A face touching at most PAIRS_PER_FACE (16) tiles is listed in each of them; otherwise it is
// listed in the coarse cells (4 x 4 tiles) it touches if those are at most 16; otherwise it goes to a per-view
// "large" list.  A tile's consumer reads its own list as is and FILTERS the enclosing coarse cell's list and
// the large list against the tile rectangle (ballot + prefix-scan compaction).  Every face occupies at most
// 16 slots at exactly one level, so the storage is bounded by PAIRS_PER_FACE * faces and cannot overflow.  Lists are UNORDERED; consumers that need
// face order (depth ties, soft-mask "first knum faces") use order-independent rules or sort in smem.
#pragma once
#include "dibr_common.cuh"
#include "dibr_prepare.cuh"      // project_face: the mesh form of setup_count projects the faces itself

namespace dibr {

struct BinLists {
  uint2* ranges;      // [Ntot]  packed inclusive pixel range per face (empty_range() = contributes nothing)
  int* tile_count;    // [B*S]   S = slots_per_view: per view, the fine tiles then the coarse cells
  int* tile_offset;   // [B*S]
  int* tile_cursor;   // [B*S]   fill cursor (absolute position)
  int* pairs;         // [PAIRS_PER_FACE * Ntot] face ids local to their view
  int* large_list;    // [Ntot]  view b's entries start at view_base(b)
  int* large_count;   // [B]
  int* alloc_cursor;  // [1]
  int* tile_order;    // [B*tiles_per_view] per view: its tiles sorted by decreasing candidate count (log2 buckets)
  int* bucket_hist;   // [B*NBUCKET] tiles per weight bucket (zeroed with the counters)
  int* bucket_cursor; // [B*NBUCKET] placement cursors of fill's ordering step (zeroed with the counters)
};
constexpr int NBUCKET = 33;            // floor(log2(weight)) + 1, weight < 2^32
struct BinLists2 { BinLists l[2]; };   // [0] rasterizer, [1] soft mask (or one set twice)

constexpr int MAX_FEATS = 8;
template <typename T>
struct FeatSet {      // up to MAX_FEATS feature tensors with the same leading dims, C_t channels each
  T* ptr[MAX_FEATS];
  int ch[MAX_FEATS];
  int n;
};

struct FaceLayout {   // batched (B,F) layout or packed layout with first_idx (B+1) on device
  const int64_t* first_idx;   // nullptr => batched
  int F;
  int64_t Ntot;
  int B;
  __device__ __forceinline__ int64_t base(int b) const { return first_idx ? first_idx[b] : (int64_t)b * F; }
  __device__ __forceinline__ int view_of(int64_t g) const {
    if (!first_idx) return (int)((unsigned)g / (unsigned)F);     // Ntot < 2^31 (check_dims): 32-bit division
    int lo = 0, hi = B;                // largest b with first_idx[b] <= g
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (first_idx[mid] <= g) lo = mid; else hi = mid;
    }
    return lo;
  }
};

template <typename T>
struct SetupParams {
  const T* img;            // (Ntot,3,2)
  const T* bbox_in;        // (Ntot,4) already multiplied/enlarged, or nullptr => computed here
  const uint8_t* valid;    // (Ntot) or nullptr
  const T* normals_z;      // (Ntot) or nullptr: valid = normals_z >= 0 (dibr_rasterization, dibr.py:195)
  T coord_mult;            // img is multiplied by this (1 when already multiplied)
  int premultiplied;
  T margin;                // bbox enlargement in multiplied units (0 for the rasterizer)
  float pix_mult;          // the kernel-side `float multiplier` of the pixel centres
};

template <typename T>
__device__ __forceinline__ void load_face_xy(const T* img, int64_t g, T coord_mult, int premultiplied, T* p) {
  const T* s = img + g * 6;
#pragma unroll
  for (int k = 0; k < 6; ++k) p[k] = premultiplied ? s[k] : RN<T>::mul(s[k], coord_mult);
}
template <>
__device__ __forceinline__ void load_face_xy<float>(const float* img, int64_t g, float coord_mult,
                                                    int premultiplied, float* p) {
  const float2* s = reinterpret_cast<const float2*>(img + g * 6);     // 24-byte records: 8-byte aligned
  const float2 a = __ldg(s), b = __ldg(s + 1), c = __ldg(s + 2);
  p[0] = a.x; p[1] = a.y; p[2] = b.x; p[3] = b.y; p[4] = c.x; p[5] = c.y;
  if (!premultiplied) {
#pragma unroll
    for (int k = 0; k < 6; ++k) p[k] = __fmul_rn(p[k], coord_mult);
  }
}

// Counting (FILL = false: K0) or scattering (FILL = true: K1c) of one face per lane into the tile lists; called by
// whole warps.  Consecutive faces of a mesh touch the same few tiles, so their atomics hit the same counters
// (ncu: the enlarged soft-mask boxes of config 3 put ~900 atomics on every counter and fill_kernel waited 150
// cycles per issued instruction).  When the fine-level faces of the warp lie in one view and the union of their tile
// rectangles has at most 64 tiles, the warp walks the union: one ballot per tile, ONE atomic per (warp, tile) by
// lane 0, positions by popc rank (face order inside the warp).  Otherwise (random soups, view boundaries) and for
// the coarse level every lane issues its own atomics, as before.  `has` = the lane holds a face with a non-empty
// pixel range; (tx0..tx1, ty0..ty1) its tile rectangle; `local` its index in view `b`.
// One level (fine tiles or coarse cells) of bin_faces_warp: lanes with `act` hold the cell rectangle (x0..x1, y0..y1)
// of their face in a grid that is `stride` cells wide, whose counters start at `slot0` of the view's slots.
// 32 x 32 bit transpose across the warp: lane i passes row i and gets column i (bit r of the result = bit i of
// lane r's argument).  Five butterfly stages, one shuffle each.
__device__ __forceinline__ unsigned warp_transpose32(unsigned a) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 16; j > 0; j >>= 1) {
    const unsigned m = j == 16 ? 0x0000ffffu : j == 8 ? 0x00ff00ffu : j == 4 ? 0x0f0f0f0fu : j == 2 ? 0x33333333u
                                                                                                    : 0x55555555u;
    const unsigned o = __shfl_xor_sync(0xffffffffu, a, j);
    a = (lane & j) ? ((a & ~m) | ((o & ~m) >> j)) : ((a & m) | ((o & m) << j));
  }
  return a;
}

template <bool FILL>
__device__ __forceinline__ void bin_level_warp(const bool act, const int b, const int local, const int x0,
                                               const int x1, const int y0, const int y1, const int slot0,
                                               const int stride, const Geometry& geo, const BinLists& L) {
  const int lane = threadIdx.x & 31;
  const unsigned am = __ballot_sync(0xffffffffu, act);
  if (am == 0u) return;
  int* const ctr0 = (FILL ? L.tile_cursor : L.tile_count) + slot0;
  const int b0 = __shfl_sync(0xffffffffu, b, __ffs(am) - 1);
  const bool same = __all_sync(0xffffffffu, !act || b == b0);
  const int ux0 = __reduce_min_sync(0xffffffffu, act ? x0 : 0x7fffffff);
  const int uy0 = __reduce_min_sync(0xffffffffu, act ? y0 : 0x7fffffff);
  const int ux1 = __reduce_max_sync(0xffffffffu, act ? x1 : -1);
  const int uy1 = __reduce_max_sync(0xffffffffu, act ? y1 : -1);
  const int uw = ux1 - ux0 + 1, ncell = uw * (uy1 - uy0 + 1);
  if (same && ncell <= 64) {
    // The warp's (face, cell) incidences as a 32 x 64 bit matrix: lane = face holds the row-major mask of the union
    // cells its rectangle covers (<= 16 bits); one bit-matrix transpose (two when the union has more than 32 cells)
    // hands lane c the set of faces in union cells c and c + 32 — their popcounts are the cells' counts, and a face's
    // position inside a cell is its rank in that set (face order inside the warp).  Lane c issues the ONE atomic of
    // (warp, cell c); no walk over the union's cells (that walk — a ballot and four compares per cell — was 50 % of
    // setup_count's and 85 % of fill's instructions: ncu, profiles/r02_a_c3_b8_ncu_summary.csv).
    int* ctr = ctr0 + (int64_t)b0 * geo.slots_per_view;
    unsigned long long cells = 0ull;
    if (act) {
      const unsigned long long row = ((1ull << (x1 - x0 + 1)) - 1ull) << (x0 - ux0);
      for (int y = y0; y <= y1; ++y) cells |= row << ((y - uy0) * uw);
    }
    const bool two = ncell > 32;                                      // warp-uniform
    const unsigned t_lo = warp_transpose32((unsigned)cells);
    const unsigned t_hi = two ? warp_transpose32((unsigned)(cells >> 32)) : 0u;
    // row / column of union cells `lane` and `lane + 32`: exact for c < 64, uw <= 64 (the quotient's fraction is at
    // least 1 / 128 away from an integer, the product's error below 1e-5)
    const float inv_uw = __frcp_rn((float)uw);
    const int r_lo = (int)(((float)lane + 0.5f) * inv_uw), r_hi = (int)(((float)(lane + 32) + 0.5f) * inv_uw);
    int* const p_lo = ctr + (uy0 + r_lo) * stride + ux0 + (lane - r_lo * uw);
    int* const p_hi = ctr + (uy0 + r_hi) * stride + ux0 + (lane + 32 - r_hi * uw);
    if (!FILL) {
      if (t_lo != 0u) atomicAdd(p_lo, __popc(t_lo));                  // no return value: a RED
      if (t_hi != 0u) atomicAdd(p_hi, __popc(t_hi));
      return;
    }
    // FILL: all of the warp's reservations are in flight together (one round trip); then every face walks its OWN
    // cells (<= 16) and takes the cell's base and member set from the lane that owns the cell.
    int base_lo = 0, base_hi = 0;
    if (t_lo != 0u) base_lo = atomicAdd(p_lo, __popc(t_lo));
    if (t_hi != 0u) base_hi = atomicAdd(p_hi, __popc(t_hi));
    const int n_own = act ? (x1 - x0 + 1) * (y1 - y0 + 1) : 0;
    const int n_max = __reduce_max_sync(0xffffffffu, n_own);
    const unsigned lt = (1u << lane) - 1u;
    int cx = x0, cy = y0;
    for (int k = 0; k < n_max; ++k) {                                 // warp-uniform trip count
      const bool mine = k < n_own;
      const int c = mine ? (cy - uy0) * uw + (cx - ux0) : 0;
      unsigned t = __shfl_sync(0xffffffffu, t_lo, c & 31);
      int base = __shfl_sync(0xffffffffu, base_lo, c & 31);
      if (two) {
        const unsigned t2 = __shfl_sync(0xffffffffu, t_hi, c & 31);
        const int base2 = __shfl_sync(0xffffffffu, base_hi, c & 31);
        if (c >= 32) { t = t2; base = base2; }
      }
      if (mine) {
        L.pairs[base + __popc(t & lt)] = local;
        if (++cx > x1) { cx = x0; ++cy; }
      }
    }
    return;
  }
  if (!act) return;
  int* ctr = ctr0 + (int64_t)b * geo.slots_per_view;
  for (int y = y0; y <= y1; ++y)
    for (int x = x0; x <= x1; ++x) {
      const int pos = atomicAdd(ctr + y * stride + x, 1);
      if (FILL) L.pairs[pos] = local;
    }
}

// Counting (FILL = false: K0) or scattering (FILL = true: K1c) of one face per lane into the lists; called by whole
// warps.  Consecutive faces of a mesh touch the same few cells, so their atomics hit the same counters (ncu: the
// enlarged soft-mask boxes of config 3 put ~900 atomics on every tile counter and fill_kernel waited 150 cycles per
// issued instruction; at config 5 the pole slivers put tens of thousands on a few coarse cells: 9.7 of the step's
// 18 ms).  At each level, when the warp's faces lie in one view and the union of their cell rectangles has at most
// 64 cells, the warp walks the union: one ballot per cell, ONE atomic per (warp, cell) by lane 0, positions by popc
// rank (face order inside the warp).  Otherwise (random soups, view boundaries) every lane issues its own atomics.
// `has` = the lane holds a face with a non-empty pixel range; (tx0..tx1, ty0..ty1) its tile rectangle; `local` its
// index in view `b`.
template <bool FILL>
__device__ __forceinline__ void bin_faces_warp(const bool has, const int b, const int local, const int tx0,
                                               const int tx1, const int ty0, const int ty1, const int64_t g,
                                               const FaceLayout& fl, const Geometry& geo, const BinLists& L) {
  const int lane = threadIdx.x & 31;
  const bool fine = has && (tx1 - tx0 + 1) * (ty1 - ty0 + 1) <= PAIRS_PER_FACE;
  bin_level_warp<FILL>(fine, b, local, tx0, tx1, ty0, ty1, 0, geo.tiles_x, geo, L);
  const int cx0 = tx0 / COARSE, cx1 = tx1 / COARSE, cy0 = ty0 / COARSE, cy1 = ty1 / COARSE;
  const bool not_fine = has && !fine;
  const bool coarse = not_fine && (cx1 - cx0 + 1) * (cy1 - cy0 + 1) <= PAIRS_PER_FACE;
  if (__any_sync(0xffffffffu, not_fine)) {
    bin_level_warp<FILL>(coarse, b, local, cx0, cx1, cy0, cy1, geo.tiles_per_view, geo.coarse_x, geo, L);
    if (!FILL) {                                         // per-view large list: written once, by the counting pass
      const bool large = not_fine && !coarse;
      const unsigned lm = __ballot_sync(0xffffffffu, large);
      if (lm != 0u) {
        const int b0 = __shfl_sync(0xffffffffu, b, __ffs(lm) - 1);
        if (__all_sync(0xffffffffu, !large || b == b0)) {  // one reservation for the warp
          int base = 0;
          if (lane == 0) base = atomicAdd(L.large_count + b0, __popc(lm));
          base = __shfl_sync(0xffffffffu, base, 0);
          if (large) L.large_list[fl.base(b0) + base + __popc(lm & ((1u << lane) - 1u))] = local;
        } else if (large) {
          L.large_list[fl.base(b) + atomicAdd(L.large_count + b, 1)] = local;
        }
      }
    }
  }
}

// bbox (multiplied coordinates) -> exact inclusive pixel range of the reference's half-open test, record + tile rect
template <typename T>
__device__ __forceinline__ bool face_pixel_range(const T xmin, const T ymin, const T xmax, const T ymax,
                                                 const float pix_mult, const Geometry& geo, uint2& rec, int& tx0,
                                                 int& tx1, int& ty0, int& ty1) {
  const float sx = pix_scale(pix_mult, geo.W), sy = pix_scale(pix_mult, geo.H);
  int ix0, ix1, ky0, ky1;
  pass_range<T>(sx, geo.W, xmin, xmax, ix0, ix1);
  pass_range<T>(sy, geo.H, ymin, ymax, ky0, ky1);      // k counts rows from the bottom
  if (ix0 > ix1 || ky0 > ky1) return false;
  const int jy0 = geo.H - 1 - ky1, jy1 = geo.H - 1 - ky0;
  rec = pack_range(ix0, ix1, jy0, jy1);
  tx0 = ix0 / TILE; tx1 = ix1 / TILE; ty0 = jy0 / TILE; ty1 = jy1 / TILE;
  return true;
}

// DUAL = false: one set of lists (L.l[0]) with sp.margin / sp.valid as given.  DUAL = true: L.l[0] = rasterizer lists
// (margin 0, only faces with valid != 0 / normals_z >= 0), L.l[1] = soft-mask lists (margin soft_margin, ALL
// faces: dibr.py:75-117 takes no valid mask), both from one read of the coordinates.
// MESH (with DUAL; dibr_rasterization_from_mesh): the thread projects its face itself from (vertices, faces, camera)
// with prepare_vertices' arithmetic (project_face) — prepare_vertices (utils.py:128-175) folded into K0 — writes
// face_vertices_image (B,F,3,2) and face_vertices_z (B,F,3) for the per-tile kernels and the backward, and culls with
// the unit normal's z (dibr.py:195); neither the camera-space (B,F,3,3) tensor nor the normals are materialised.
template <typename T>
struct MeshOut {
  T* img;      // (B,F,3,2)
  T* z;        // (B,F,3)
};
template <typename T, bool DUAL, bool MESH>
__global__ void __launch_bounds__(256) setup_count_kernel(SetupParams<T> sp, const T soft_margin, FaceLayout fl,
                                                          Geometry geo, BinLists2 L, PrepareParams<T> mp,
                                                          MeshOut<T> mo) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool in_range = g < fl.Ntot;                    // no early exit: the binning below is warp-collective
  bool is_valid = in_range;
  if (!MESH) {
    if (in_range && sp.valid) is_valid = sp.valid[g] != 0;
    if (in_range && sp.normals_z) is_valid = sp.normals_z[g] >= (T)0;   // dibr.py:195 `face_normals_z >= 0.`
  }
  uint2 rec = empty_range(), rec2 = empty_range();
  int b = 0, tx0 = 0, tx1 = -1, ty0 = 0, ty1 = -1;
  int sx0 = 0, sx1 = -1, sy0 = 0, sy1 = -1;
  bool has = false, has2 = false;
  if (MESH || DUAL ? in_range : is_valid) {
    T xmin, ymin, xmax, ymax;
    if (MESH) {
      b = fl.view_of(g);
      const ViewCamera<T> cam = load_camera<T>(mp, b);
      T q[3][3], im[6], nrm[3], p[6];
      project_face<T>(mp, cam, b, (int)(g - fl.base(b)), q, im, nrm);
      is_valid = nrm[2] >= (T)0;
      T* oi = mo.img + g * 6;
      T* oz = mo.z + g * 3;
#pragma unroll
      for (int k = 0; k < 6; ++k) { oi[k] = im[k]; p[k] = RN<T>::mul(im[k], sp.coord_mult); }
      oz[0] = q[0][2]; oz[1] = q[1][2]; oz[2] = q[2][2];
      xmin = min_nan(min_nan(p[0], p[2]), p[4]);
      xmax = max_nan(max_nan(p[0], p[2]), p[4]);
      ymin = min_nan(min_nan(p[1], p[3]), p[5]);
      ymax = max_nan(max_nan(p[1], p[3]), p[5]);
    } else if (sp.bbox_in) {
      const T* bb = sp.bbox_in + g * 4;
      xmin = bb[0]; ymin = bb[1]; xmax = bb[2]; ymax = bb[3];
    } else {
      T p[6];
      load_face_xy<T>(sp.img, g, sp.coord_mult, sp.premultiplied, p);
      xmin = min_nan(min_nan(p[0], p[2]), p[4]);
      xmax = max_nan(max_nan(p[0], p[2]), p[4]);
      ymin = min_nan(min_nan(p[1], p[3]), p[5]);
      ymax = max_nan(max_nan(p[1], p[3]), p[5]);
    }
    if (DUAL) {
      if (is_valid) has = face_pixel_range<T>(xmin, ymin, xmax, ymax, sp.pix_mult, geo, rec, tx0, tx1, ty0, ty1);
      // dibr.py:36-38: points_min - boxlen*multiplier, points_max + boxlen*multiplier
      has2 = face_pixel_range<T>(RN<T>::sub(xmin, soft_margin), RN<T>::sub(ymin, soft_margin),
                                 RN<T>::add(xmax, soft_margin), RN<T>::add(ymax, soft_margin), sp.pix_mult, geo,
                                 rec2, sx0, sx1, sy0, sy1);
    } else {
      if (!sp.bbox_in) {
        xmin = RN<T>::sub(xmin, sp.margin); ymin = RN<T>::sub(ymin, sp.margin);
        xmax = RN<T>::add(xmax, sp.margin); ymax = RN<T>::add(ymax, sp.margin);
      }
      has = face_pixel_range<T>(xmin, ymin, xmax, ymax, sp.pix_mult, geo, rec, tx0, tx1, ty0, ty1);
    }
    if (!MESH && (has || has2)) b = fl.view_of(g);
  }
  const int local = (has || has2) ? (int)(g - fl.base(b)) : 0;
  bin_faces_warp<false>(has, b, local, tx0, tx1, ty0, ty1, g, fl, geo, L.l[0]);
  if (in_range) L.l[0].ranges[g] = rec;
  if (DUAL) {
    bin_faces_warp<false>(has2, b, local, sx0, sx1, sy0, sy1, g, fl, geo, L.l[1]);
    if (in_range) L.l[1].ranges[g] = rec2;
  }
}

// scheduling weight of a tile: its own list + a share of the enclosing coarse cell's list (filtered by the consumer)
__device__ __forceinline__ int tile_weight_bucket(const int* cnt, const Geometry& geo, const int t) {
  const int tx = t % geo.tiles_x, ty = t / geo.tiles_x;
  const int w = cnt[t] + (cnt[geo.tiles_per_view + (ty / COARSE) * geo.coarse_x + (tx / COARSE)] >> 2);
  return w > 0 ? 32 - __clz(w) : 0;
}

// grid (chunks of SCAN_CHUNK counters, views, list sets): see the file header (K1b)
constexpr int SCAN_CHUNK = 1024;
__global__ void __launch_bounds__(SCAN_CHUNK) scan_kernel(Geometry geo, BinLists2 LL) {
  __shared__ int warp_sums[32];
  __shared__ int s_base;
  __shared__ int hist[NBUCKET];
  const BinLists& L = LL.l[blockIdx.z];
  const int b = blockIdx.y;
  const int T = geo.slots_per_view;
  const int* cnt = L.tile_count + (int64_t)b * T;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int t = blockIdx.x * SCAN_CHUNK + threadIdx.x;
  if (threadIdx.x < NBUCKET) hist[threadIdx.x] = 0;
  const int c = (t < T) ? cnt[t] : 0;
  int incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += n;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const int v = warp_sums[lane];
    int wi = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += n;
    }
    warp_sums[lane] = wi - v;                                     // exclusive prefix of the warp totals
    if (lane == 31) s_base = wi > 0 ? atomicAdd(L.alloc_cursor, wi) : 0;    // this chunk's segment of `pairs`
  }
  // bucket counts of the chunk's fine tiles (heaviest-first launch order of the per-tile kernels: a few tiles hold
  // hundreds of sliver faces and would otherwise finish long after the rest of the grid)
  if (t < geo.tiles_per_view) atomicAdd(&hist[tile_weight_bucket(cnt, geo, t)], 1);
  __syncthreads();
  if (t < T) {
    const int o = s_base + warp_sums[warp] + incl - c;
    L.tile_offset[(int64_t)b * T + t] = o;
    L.tile_cursor[(int64_t)b * T + t] = o;
  }
  if (threadIdx.x < NBUCKET && hist[threadIdx.x] > 0)
    atomicAdd(&L.bucket_hist[b * NBUCKET + threadIdx.x], hist[threadIdx.x]);
}
inline dim3 scan_grid(const Geometry& geo, int n_sets) {
  return dim3((unsigned)((geo.slots_per_view + SCAN_CHUNK - 1) / SCAN_CHUNK), (unsigned)geo.B, (unsigned)n_sets);
}

// Ordering step of fill: thread = (view, tile).  Its bucket's first position = the number of tiles in heavier buckets
// (33 cached loads); lanes of a warp that share (view, bucket) — 32 neighbouring tiles usually do — reserve their
// positions with one atomic.  Entries are packed tile coordinates (tx | ty << 16); the order inside a bucket is
// arbitrary (it only affects scheduling).
__device__ __forceinline__ void place_tile_order(const Geometry& geo, const BinLists& L, const int64_t i) {
  const int64_t n = (int64_t)geo.B * geo.tiles_per_view;
  const bool act = i < n;
  const int b = act ? (int)(i / geo.tiles_per_view) : 0;
  const int t = act ? (int)(i - (int64_t)b * geo.tiles_per_view) : 0;
  int key = -1 - (int)(threadIdx.x & 31);
  int k = 0;
  if (act) {
    k = tile_weight_bucket(L.tile_count + (int64_t)b * geo.slots_per_view, geo, t);
    key = b * NBUCKET + k;
  }
  const unsigned peers = __match_any_sync(0xffffffffu, key);
  if (!act) return;
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(peers) - 1;
  int pos = 0;
  if (lane == leader) {
    int base = 0;
    for (int h = k + 1; h < NBUCKET; ++h) base += __ldg(L.bucket_hist + b * NBUCKET + h);
    pos = base + atomicAdd(&L.bucket_cursor[key], __popc(peers));
  }
  pos = __shfl_sync(peers, pos, leader) + __popc(peers & ((1u << lane) - 1u));
  L.tile_order[(int64_t)b * geo.tiles_per_view + pos] = (t % geo.tiles_x) | ((t / geo.tiles_x) << 16);
}

// Grid of the per-tile kernels: x = view, (y, z) = rank of the tile in its view's order — consecutive CTAs take the
// same rank in different views, ranks go from the heaviest tile of each view to the lightest.  No integer division
// in the consumers: the order entries are packed tile coordinates.
constexpr unsigned SCHED_Y = 32768u;
inline dim3 scheduled_grid(const Geometry& geo) {
  const unsigned t = (unsigned)geo.tiles_per_view;
  const unsigned y = t < SCHED_Y ? t : SCHED_Y;
  return dim3((unsigned)geo.B, y, (t + y - 1) / (y ? y : 1));
}
__device__ __forceinline__ bool scheduled_tile(const BinLists& L, const Geometry& geo, int& b, int& tx, int& ty) {
  const unsigned rank = blockIdx.y + blockIdx.z * gridDim.y;
  if (rank >= (unsigned)geo.tiles_per_view) return false;
  b = (int)blockIdx.x;
  const unsigned e = (unsigned)L.tile_order[(int64_t)b * geo.tiles_per_view + rank];
  tx = (int)(e & 0xffffu); ty = (int)(e >> 16);
  return true;
}

// grid covers max(Ntot, B * tiles_per_view) threads: thread = face (scatter) and thread = (view, tile) (order)
template <bool DUAL>
__global__ void __launch_bounds__(256) fill_kernel(FaceLayout fl, Geometry geo, BinLists2 LL) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
  for (int set = 0; set < (DUAL ? 2 : 1); ++set) {
    const BinLists& L = LL.l[set];
    place_tile_order(geo, L, g);
    if ((int64_t)blockIdx.x * blockDim.x >= fl.Ntot) continue;      // CTA-uniform: this CTA holds no faces
    bool has = false;
    int b = 0, local = 0, tx0 = 0, tx1 = -1, ty0 = 0, ty1 = -1;
    if (g < fl.Ntot) {
      const uint2 r = L.ranges[g];
      if (!range_empty(r)) {
        has = true;
        tx0 = (int)(r.x & 0xffffu) / TILE; tx1 = (int)(r.x >> 16) / TILE;
        ty0 = (int)(r.y & 0xffffu) / TILE; ty1 = (int)(r.y >> 16) / TILE;
        b = fl.view_of(g);
        local = (int)(g - fl.base(b));
      }
    }
    bin_faces_warp<true>(has, b, local, tx0, tx1, ty0, ty1, g, fl, geo, L);
  }
}
inline unsigned binning_blocks(const FaceLayout& fl, const Geometry& geo, bool with_order) {
  int64_t n = fl.Ntot;
  const int64_t tiles = (int64_t)geo.B * geo.tiles_per_view;
  if (with_order && tiles > n) n = tiles;
  return (unsigned)((n + 255) / 256);
}

// The candidate sources of one 16x16 tile: the faces binned to the tile itself (all touch it), those binned to
// the enclosing coarse cell and the view's "large" list (both must be filtered with range_hits_rect).
struct TileLists {
  const int* fine; int n_fine;
  const int* coarse; int n_coarse;
  const int* large; int n_large;
  __device__ __forceinline__ int n_filtered() const { return n_coarse + n_large; }
  __device__ __forceinline__ int filtered_at(int k) const { return k < n_coarse ? coarse[k] : large[k - n_coarse]; }
};
__device__ __forceinline__ TileLists tile_lists(const BinLists& L, const Geometry& geo, const FaceLayout& fl,
                                                const int b, const int tx, const int ty) {
  TileLists t;
  const int64_t s0 = (int64_t)b * geo.slots_per_view;
  const int64_t sf = s0 + ty * geo.tiles_x + tx;
  const int64_t sc = s0 + geo.tiles_per_view + (ty / COARSE) * geo.coarse_x + (tx / COARSE);
  t.fine = L.pairs + L.tile_offset[sf];     t.n_fine = L.tile_count[sf];
  t.coarse = L.pairs + L.tile_offset[sc];   t.n_coarse = L.tile_count[sc];
  t.large = L.large_list + fl.base(b);      t.n_large = L.large_count[b];
  return t;
}

// ---------------------------------------------------------------- block-level ordered compaction
// Every thread of a 256-thread CTA passes `flag`; returns the thread's slot among the flagged threads
// (in thread order) and the block total.  `scratch` = 9 ints of shared memory.
__device__ __forceinline__ int block_compact_256(bool flag, int* scratch, int& total) {
  const unsigned m = __ballot_sync(0xffffffffu, flag);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int in_warp = __popc(m & ((1u << lane) - 1u));
  __syncthreads();
  if (lane == 0) scratch[warp] = __popc(m);
  __syncthreads();
  int pre = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    const int c = scratch[w];
    if (w < warp) pre += c;
    tot += c;
  }
  total = tot;
  return pre + in_warp;
}

}  // namespace dibr
