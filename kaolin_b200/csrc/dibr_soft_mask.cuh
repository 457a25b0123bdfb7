// dibr_soft_mask.cuh — K3/K5: DIB-R soft mask forward and backward (reference:
// dibr_soft_mask_cuda.cu:27-184 and :230-353).
//
// One CTA = one 16x16 tile.  Only tiles with at least one uncovered pixel do any work beyond the
// face_idx read / soft write.  The reference keeps the FIRST `knum` faces in face order whose enlarged bbox
// contains the pixel (:80-171), so order is semantic: the tile's candidate faces (enlarged-bbox tile list +
// filtered coarse / "large" faces) are taken in passes of increasing face-index ranges — a pass holds the lowest
// remaining ids that fit the shared-memory arrays (1024 / 2048), chosen by one histogram sweep (select_hi) — gathered
// into shared memory, sorted by face index (bucket_sort / chunk_rank_sort) and turned into coverage masks of the tile's columns and rows, stored transposed per
// chunk of 32 candidates (one word per column / row = the chunk's candidates covering it).
//
// Per pass: (A, once) compact the uncovered pixels (from the rasterizer's coverage map when it left one); (B) every
// pixel gets the slots of its next hits, in face order, until it has knum — lane = pixel, one AND of its column and
// row words tests 32 candidates; (C) the (pixel, k) pairs of a warp's pixels are flattened and evaluated one per
// lane; (D) one lane per pixel multiplies the probabilities in face order.
//
// MODE_FWD writes soft (+ optional k-list side buffers, API-visible through
// _C.render.mesh.dibr_soft_mask_forward_cuda).  MODE_BWD re-derives the same k-list with the same
// arithmetic instead of reading 13*knum bytes per pixel of side buffers, and accumulates the vertex
// gradients with red.global.add.
#pragma once
#include "dibr_binning.cuh"
#include "dibr_raster_bwd.cuh"      // group_partners / group_sum (warp-aggregated atomics)

namespace dibr {

#ifndef DIBR_SOFT_TPS
#define DIBR_SOFT_TPS 1024       // resident threads per SM the fp32 register allocation aims at
#endif
#ifndef DIBR_SOFT_CELLS
#define DIBR_SOFT_CELLS 32       // k-list cells per thread of the CTA
#endif
constexpr int SOFT_CAP = 2048;      // candidate faces per tile held in shared memory
constexpr int MODE_FWD = 0;
constexpr int MODE_BWD = 1;

template <typename T>
struct SoftParams {
  const T* img;               // (B,F,3,2)
  T coord_mult;
  int premultiplied;
  float pix_mult;             // kernel-side float multiplier
  float sigmainv;
  int F, knum;
  const int64_t* face_idx;    // (B,H,W)
  // dibr_rasterization (fp32): per tile the 256-bit map of uncovered pixels left by the rasterizer's epilogue, which
  // has also written soft = 1 / 0 for every pixel (dibr_raster_fwd.cuh: RasterFwdParams::tile_unc); nullptr = derive
  // the map from face_idx here and write the covered pixels' 1
  const unsigned* tile_unc;   // [B * tiles_per_view * 8] or nullptr
  // forward
  T* soft;                    // (B,H,W)
  T* prob;                    // (B,H,W,K) or nullptr
  int64_t* close_idx;         // (B,H,W,K) or nullptr
  uint8_t* dist_type;         // (B,H,W,K) or nullptr
  // backward
  const T* grad_soft;         // (B,H,W)
  const T* soft_in;           // (B,H,W)
  T* grad_img;                // (B,F,3,2)
  // pair log (optional; written by the fused forward, consumed by soft_mask_bwd_log_kernel): every evaluated
  // (pixel, face) pair as one 16-byte record {global pixel index, (view * F + face) << 3 | distance type, probability}, in LOG_SEGMENTS
  // independent segments (a tile appends to segment hash(tile), one atomicAdd per warp and super-batch).
  // A tile whose reservation does not fit its segment raises its flag and enters the overflow list: its logged
  // pairs are ignored and soft_mask_overflow_kernel recomputes that tile (MODE_BWD).
  unsigned* log_count;        // [LOG_SEGMENTS] or nullptr = no log
  unsigned* over_count;       // [1] number of overflowed tiles
  unsigned* tile_flag;        // [B*tiles] 1 = overflowed
  unsigned* over_list;        // [B*tiles] the overflowed tiles (tile_global), in no particular order
  uint4* log;                 // [LOG_SEGMENTS * log_cap] 16-byte records {pixel, face << 3 | type, prob bits (z[, w])}
  unsigned log_cap;           // entries per segment
  int sel_first;              // candidates taken by the first pass of a tile (0 = default 2 * threads per CTA)
};
// fp32 records carry the pixel's column | row << 16 in their fourth word (H, W <= 32768), which saves the
// backward two integer divisions per pair; fp64 records need that word for the probability.
__device__ __forceinline__ uint4 log_pack(unsigned pix, unsigned fk, float prob, unsigned pxpy) {
  return make_uint4(pix, fk, __float_as_uint(prob), pxpy);
}
__device__ __forceinline__ uint4 log_pack(unsigned pix, unsigned fk, double prob, unsigned) {
  const unsigned long long u = (unsigned long long)__double_as_longlong(prob);
  return make_uint4(pix, fk, (unsigned)(u & 0xffffffffull), (unsigned)(u >> 32));
}
template <typename T> __device__ __forceinline__ T log_prob(const uint4& r);
template <> __device__ __forceinline__ float log_prob<float>(const uint4& r) { return __uint_as_float(r.z); }
template <> __device__ __forceinline__ double log_prob<double>(const uint4& r) {
  return __longlong_as_double((long long)(((unsigned long long)r.w << 32) | r.z));
}
constexpr int LOG_SEGMENTS = 64;
#ifndef DIBR_LOG_CTAS_PER_SEGMENT
#define DIBR_LOG_CTAS_PER_SEGMENT 64   // measured 16 / 32 / 64 / 128 at c3-64: 0.564 / 0.478 / 0.447 / 0.443 ms (c3-8: 0.083 / 0.073 / 0.071 / 0.075)
#endif
constexpr int LOG_CTAS_PER_SEGMENT = DIBR_LOG_CTAS_PER_SEGMENT;

// dibr_soft_mask_cuda.cu:283-347 for one (pixel, face) pair; p = multiplied coordinates.
// Gradients need 1e-4 relative, not bit-exactness (the reference accumulates them with unordered atomics): the
// divisions by `multiplier` and by the edge's squared length become multiplications by one reciprocal each —
// besides being shorter, this avoids the slow path of the IEEE division, which the many denormal-sized terms
// (dLdz ~ 1e-40 far from an edge) kept hitting (ncu: 10 % of the kernel's instructions).
template <typename T>
__device__ __forceinline__ void soft_backward_pair(const T* p, const T x0, const T y0, const T prob,
                                                   const int edgeid, const T dLdp, const T allprob,
                                                   const float sigmainv, const float multiplier, T* gi) {
  const double EPS = 1e-7;
  const T dLdz = (T)(-1.0 * (double)sigmainv * (double)dLdp * (1.0 - (double)allprob) /
                     (1.0 - (double)prob + EPS) * (double)prob);
  if (dLdz == (T)0) return;                              // nothing to add (prob or upstream gradient is zero)
  const T s = dLdz * ((T)1 / (T)multiplier);
  // selects instead of p[edgeid-dependent index]: keeps p[] in registers (no local-memory frame)
  if (edgeid >= 3) {
    const int v = edgeid - 3;
    const T x1 = v == 0 ? p[0] : (v == 1 ? p[2] : p[4]), y1 = v == 0 ? p[1] : (v == 1 ? p[3] : p[5]);
    atomicAdd(gi + 2 * v, s * 2 * (x1 - x0));
    atomicAdd(gi + 2 * v + 1, s * 2 * (y1 - y0));
  } else {
    const int e = edgeid, e2 = edgeid == 2 ? 0 : edgeid + 1;
    const T x1 = e == 0 ? p[0] : (e == 1 ? p[2] : p[4]), y1 = e == 0 ? p[1] : (e == 1 ? p[3] : p[5]);
    const T x2 = e2 == 0 ? p[0] : (e2 == 1 ? p[2] : p[4]), y2 = e2 == 0 ? p[1] : (e2 == 1 ? p[3] : p[5]);
    const T A = y2 - y1, Bc = x1 - x2, Cc = x2 * y1 - x1 * y2;
    const T up = A * x0 + Bc * y0 + Cc;
    const T down = A * A + Bc * Bc;
    const double rden = 1.0 / ((double)down + EPS);
    const T dissquare = (T)((double)(up * up) * rden);
    const T dzdA = (T)((double)(2 * (x0 * up - dissquare * A)) * rden);
    const T dzdB = (T)((double)(2 * (y0 * up - dissquare * Bc)) * rden);
    const T dzdC = (T)((double)(2 * up) * rden);
    atomicAdd(gi + 2 * e, s * (dzdB - y2 * dzdC));
    atomicAdd(gi + 2 * e + 1, s * (x2 * dzdC - dzdA));
    atomicAdd(gi + 2 * e2, s * (y1 * dzdC - dzdB));
    atomicAdd(gi + 2 * e2 + 1, s * (dzdA - x1 * dzdC));
  }
}

// fp32 instantiation: the same formulas in fp32 throughout (the double detour of `EPS 1e-7` costs a double division
// and a dozen conversions per pair; 1 - prob is exact in fp32 for prob >= 0.5 and the gradients are judged at 1e-4).
template <>
__device__ __forceinline__ void soft_backward_pair<float>(const float* p, const float x0, const float y0,
                                                          const float prob, const int edgeid, const float dLdp,
                                                          const float allprob, const float sigmainv,
                                                          const float multiplier, float* gi) {
  const float dLdz = -sigmainv * dLdp * (1.0f - allprob) * __fdividef(prob, (1.0f - prob) + 1e-7f);
  if (dLdz == 0.0f) return;
  const float s = dLdz * __fdividef(1.0f, multiplier);
  // a vertex's (x, y) gradient goes out as ONE 8-byte vector reduction (red.global.add.v2.f32: the (3,2) record of a
  // face is 24 bytes, every vertex 8-byte aligned) — half the atomic requests of a kernel that ncu shows throttled on
  // its load / store queue (lg_throttle)
  if (edgeid >= 3) {
    const int v = edgeid - 3;
    const float x1 = v == 0 ? p[0] : (v == 1 ? p[2] : p[4]), y1 = v == 0 ? p[1] : (v == 1 ? p[3] : p[5]);
    red_add2(gi + 2 * v, s * 2 * (x1 - x0), s * 2 * (y1 - y0));
  } else {
    const int e = edgeid, e2 = edgeid == 2 ? 0 : edgeid + 1;
    const float x1 = e == 0 ? p[0] : (e == 1 ? p[2] : p[4]), y1 = e == 0 ? p[1] : (e == 1 ? p[3] : p[5]);
    const float x2 = e2 == 0 ? p[0] : (e2 == 1 ? p[2] : p[4]), y2 = e2 == 0 ? p[1] : (e2 == 1 ? p[3] : p[5]);
    const float A = y2 - y1, Bc = x1 - x2, Cc = x2 * y1 - x1 * y2;
    const float up = A * x0 + Bc * y0 + Cc;
    const float rden = __fdividef(1.0f, (A * A + Bc * Bc) + 1e-7f);
    const float dissquare = (up * up) * rden;
    const float dzdA = 2 * (x0 * up - dissquare * A) * rden;
    const float dzdB = 2 * (y0 * up - dissquare * Bc) * rden;
    const float dzdC = 2 * up * rden;
    red_add2(gi + 2 * e, s * (dzdB - y2 * dzdC), s * (x2 * dzdC - dzdA));
    red_add2(gi + 2 * e2, s * (y1 * dzdC - dzdB), s * (dzdA - x1 * dzdC));
  }
}

// distance + probability of one (pixel, face) pair: the reference's arithmetic, or — fp32, fused path — its fp32 form
template <typename T>
__device__ __forceinline__ void soft_pair(const T* p, const T x0, const T y0, const float multiplier,
                                          const float sigmainv, const bool fused, const float prob_scale, T& d,
                                          T& prob, int& edgeid) {
  d = soft_distance<T>(p, x0, y0, multiplier, edgeid);
  prob = soft_prob<T>(d, sigmainv, multiplier);
}
template <>
__device__ __forceinline__ void soft_pair<float>(const float* p, const float x0, const float y0,
                                                 const float multiplier, const float sigmainv, const bool fused,
                                                 const float prob_scale, float& d, float& prob, int& edgeid) {
  bool ambiguous = false;
  if (fused) {
    d = soft_distance_f32(p, x0, y0, multiplier, edgeid, ambiguous);
    prob = soft_prob_f32(d, prob_scale);
  }
  if (!fused || ambiguous) {          // ties between the six distances: the reference's exact arithmetic decides
    d = soft_distance<float>(p, x0, y0, multiplier, edgeid);
    prob = soft_prob<float>(d, sigmainv, multiplier);
  }
}

template <typename T, int MODE>
struct SoftPixel {            // per-pixel running state of the k-list walk
  T x0, y0;
  int kid;
  T allprob;                  // forward: running product of (1 - prob)
  T dLdp, soft;               // backward inputs
  int64_t pix;
};

template <typename T, int MODE>
__device__ __forceinline__ void soft_visit(const SoftParams<T>& sp, SoftPixel<T, MODE>& st, const int64_t vbase,
                                           const int f) {
  T p[6];
  load_face_xy<T>(sp.img, vbase + f, sp.coord_mult, sp.premultiplied, p);
  int edgeid;
  const T d = soft_distance<T>(p, st.x0, st.y0, sp.pix_mult, edgeid);
  const T prob = soft_prob<T>(d, sp.sigmainv, sp.pix_mult);
  if (MODE == MODE_FWD) {
    if (sp.prob) {
      const int64_t o = st.pix * sp.knum + st.kid;
      sp.prob[o] = prob;
      sp.close_idx[o] = (int64_t)f;
      sp.dist_type[o] = (uint8_t)(edgeid + 1);
    }
    st.allprob = (T)((double)st.allprob * (1.0 - (double)prob));     // :174-178, same order
  } else {
    soft_backward_pair<T>(p, st.x0, st.y0, prob, edgeid, st.dLdp, st.soft, sp.sigmainv, sp.pix_mult,
                          sp.grad_img + (vbase + f) * 6);
  }
  ++st.kid;
}

// Shared-memory plan (dynamic): sorted candidate ids, their tile-local coverage masks, the compacted
// uncovered pixels and, per warp, CELLS k-list cells (a cell first holds the candidate slot of a
// (pixel, k) pair, then — forward — its probability) plus small per-batch tables.
template <typename T> struct CellInt;
template <> struct CellInt<float> { typedef unsigned type; };
template <> struct CellInt<double> { typedef unsigned long long type; };
// NT = threads per CTA (64 / 128 / 256; one CTA per tile whatever NT).  Small CTAs keep more independent tiles
// resident per SM, which matters because the per-tile phases are separated by barriers and rarely have work
// for 8 warps (ncu: barrier stalls dominated the 256-thread version).
template <typename T, int NT>
struct SoftSmem {
  static constexpr int NW = NT / 32;
  static constexpr int CAP = (NT >= 256) ? 2048 : 1024;   // candidate faces per pass
  static constexpr int CELLS_TOT = NT * DIBR_SOFT_CELLS;               // k-list cells: a super-batch is min(NT, CELLS_TOT / knum) pixels
  static constexpr size_t ids = 0;
  static constexpr size_t mask = ids + (CAP + 8) * sizeof(int);      // + padding
  static constexpr size_t cell = mask + (CAP + 8) * sizeof(unsigned);
  static constexpr size_t upix = cell + CELLS_TOT * sizeof(T);
  static constexpr size_t wtab = upix + 256 * sizeof(int);
  static constexpr size_t ktot = wtab + NW * 32 * sizeof(int);
  static constexpr size_t knew = ktot + 256 * sizeof(int);
  static constexpr size_t acc = knew + 256 * sizeof(int);
  static constexpr size_t total = acc + 256 * sizeof(T);
};

// Number of the tile's candidates (tile list + "large" faces touching the tile) with lo < id <= hi.
// CTA-wide; `scratch` = 9 ints.  Used only by the multi-pass selection of very dense tiles.
template <int NT>
__device__ __forceinline__ int count_candidates(const TileLists& tl, const uint2* ranges, const int lo,
                                                const int hi, const int rx0, const int ry0, const int rx1,
                                                const int ry1, int* scratch) {
  int c = 0;
  for (int k = threadIdx.x; k < tl.n_fine; k += NT) {
    const int f = tl.fine[k];
    c += (f > lo && f <= hi) ? 1 : 0;
  }
  const int n_filt = tl.n_filtered();
  for (int k = threadIdx.x; k < n_filt; k += NT) {
    const int f = tl.filtered_at(k);
    if (f > lo && f <= hi && range_hits_rect(ranges[f], rx0, ry0, rx1, ry1)) ++c;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = c;
  __syncthreads();
  int tot = 0;
#pragma unroll
  for (int w = 0; w < NT / 32; ++w) tot += scratch[w];
  return tot;
}

// Upper face index `hi` of the next pass: the candidates with lo < id <= hi are about `target` (at most `cap`) of
// the tile's lowest remaining ids.  One histogram sweep over the tile's lists (256 buckets of 2^shift consecutive
// ids above lo, shared-memory atomics) instead of a bisection of count_candidates() calls; `count` returns the exact
// number selected.  Returns lo when the first bucket alone holds more than `cap` candidates (the caller bisects
// inside (lo, lo + 2^shift], reported through `hi_bucket`).  CTA-wide; `hist` = 256 ints, `scratch` = 9 ints.
template <int NT>
__device__ __forceinline__ int select_hi(const TileLists& tl, const uint2* ranges, const int lo, const int nf,
                                         const int target, const int cap, const int rx0, const int ry0,
                                         const int rx1, const int ry1, int* hist, int* scratch, int& count,
                                         int& hi_bucket) {
  const int span = nf - 1 - lo;                                   // ids lo + 1 .. nf - 1
  int shift = 0;
  while (((span - 1) >> shift) > 255) ++shift;
  for (int k = threadIdx.x; k < 256; k += NT) hist[k] = 0;
  __syncthreads();
  for (int k = threadIdx.x; k < tl.n_fine; k += NT) {
    const int f = tl.fine[k];
    if (f > lo) atomicAdd(&hist[(f - lo - 1) >> shift], 1);
  }
  const int n_filt = tl.n_filtered();
  for (int k = threadIdx.x; k < n_filt; k += NT) {
    const int f = tl.filtered_at(k);
    if (f > lo && range_hits_rect(ranges[f], rx0, ry0, rx1, ry1)) atomicAdd(&hist[(f - lo - 1) >> shift], 1);
  }
  __syncthreads();
  if (threadIdx.x < 32) {                                         // warp 0: inclusive prefix over the 256 buckets
    const int lane = threadIdx.x;
    int c[8], sum = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) { c[q] = hist[lane * 8 + q]; sum += c[q]; }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    int run = incl - sum, n_cap = 0, n_tgt = 0;                    // the prefix is non-decreasing: both are counts
#pragma unroll                                                     // of a leading run of buckets
    for (int q = 0; q < 8; ++q) {
      run += c[q];
      hist[lane * 8 + q] = run;
      n_cap += run <= cap ? 1 : 0;
      n_tgt += run < target ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      n_cap += __shfl_xor_sync(0xffffffffu, n_cap, o);
      n_tgt += __shfl_xor_sync(0xffffffffu, n_tgt, o);
    }
    __syncwarp();
    // last bucket taken: the first one that reaches `target`, but never beyond the capacity
    const int t = min(min(n_cap - 1, n_tgt), 255);
    if (lane == 0) {
      scratch[0] = t; scratch[1] = t >= 0 ? hist[t] : 0; scratch[2] = shift;
    }
  }
  __syncthreads();
  const int t = scratch[0];
  count = scratch[1];
  const long long step = 1ll << scratch[2];
  hi_bucket = (int)min((long long)(nf - 1), (long long)lo + step);
  __syncthreads();                                                // scratch / hist are reused by the caller
  if (t < 0) return lo;
  return (int)min((long long)(nf - 1), (long long)lo + (long long)(t + 1) * step);
}

// Sort of `count` <= 4 * NT unique ids in shared memory (capacity: count rounded up to 32).  Chunks of 32 are
// sorted by one warp each in registers (bitonic network over shuffles); an element's final position is then its
// lane plus, for every other chunk, the number of smaller elements there (branch-free binary search over the
// sorted chunk: every word of a 32-word chunk is its own bank, so the searches are conflict-free).  About a third
// of the instructions of the all-pairs ranking it replaces (count^2 / NT compares per thread).
__device__ __forceinline__ int bitonic32(int v) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      const int o = __shfl_xor_sync(0xffffffffu, v, j);
      const bool keep_min = ((lane & j) == 0) == ((lane & k) == 0);   // (lane & 32) == 0: the last merge ascends
      v = keep_min ? min(v, o) : max(v, o);
    }
  }
  return v;
}
__device__ __forceinline__ int count_less32(const int* ch, const int x) {     // ch: 32 ascending values
  int pos = 0;
#pragma unroll
  for (int st = 16; st > 0; st >>= 1) pos += (ch[pos + st - 1] < x) ? st : 0;
  return pos + ((ch[pos] < x) ? 1 : 0);
}
template <int NT>
__device__ __forceinline__ void chunk_rank_sort(int* s_ids, const int count) {
  constexpr int NW = NT / 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nchunks = (count + 31) >> 5;                           // <= 4 * NW
  for (int k = count + threadIdx.x; k < nchunks * 32; k += NT) s_ids[k] = 0x7fffffff;
  __syncthreads();
  for (int c = warp; c < nchunks; c += NW) s_ids[c * 32 + lane] = bitonic32(s_ids[c * 32 + lane]);
  __syncthreads();
  int mine[4], rank[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int c = warp + q * NW;                                   // warp-uniform
    mine[q] = 0x7fffffff; rank[q] = 0;
    if (c < nchunks) {
      mine[q] = s_ids[c * 32 + lane];
      int r = lane;
      for (int c2 = 0; c2 < nchunks; ++c2)
        if (c2 != c) r += count_less32(s_ids + c2 * 32, mine[q]);
      rank[q] = r;
    }
  }
  __syncthreads();
#pragma unroll
  for (int q = 0; q < 4; ++q)
    if (mine[q] != 0x7fffffff) s_ids[rank[q]] = mine[q];           // ids are unique: ranks are a permutation
  __syncthreads();
}

// Sort of `count` unique ids in shared memory by distribution: 256 buckets over [min, max] of the ids (histogram with
// shared-memory atomics, one warp scans it, the ids are scattered to their buckets in `s_tmp`), then every id is
// ranked inside its bucket only.  A tile's candidates are a few runs of consecutive face ids, so a bucket holds a
// handful to a few tens of them: ~100 instructions per id instead of the ~400 of chunk_rank_sort (which ranks every
// id against every 32-chunk: 6 binary-search steps per chunk) at 256-512 candidates, and no bitonic network beyond
// 4 * NT.  `s_tmp`: count words; `s_hist`: 256 words; `scratch`: 2 words.  Worst case (all ids in one bucket): the
// all-pairs ranking of count^2 / NT compares per thread.
#ifndef DIBR_SOFT_BUCKET_SORT
#define DIBR_SOFT_BUCKET_SORT 64   // candidates per pass from which the distribution sort is used (0 = never)
#endif
template <int NT>
__device__ __forceinline__ void bucket_sort(int* s_ids, int* s_tmp, int* s_hist, int* scratch, const int count) {
  int mn = 0x7fffffff, mx = -1;
  for (int k = threadIdx.x; k < count; k += NT) { const int v = s_ids[k]; mn = min(mn, v); mx = max(mx, v); }
  mn = __reduce_min_sync(0xffffffffu, mn);
  mx = __reduce_max_sync(0xffffffffu, mx);
  for (int k = threadIdx.x; k < 256; k += NT) s_hist[k] = 0;
  if (threadIdx.x == 0) { scratch[0] = 0x7fffffff; scratch[1] = -1; }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) { atomicMin(&scratch[0], mn); atomicMax(&scratch[1], mx); }
  __syncthreads();
  mn = scratch[0]; mx = scratch[1];
  int shift = 0;
  while (((mx - mn) >> shift) > 255) ++shift;
  for (int k = threadIdx.x; k < count; k += NT) atomicAdd(&s_hist[(s_ids[k] - mn) >> shift], 1);
  __syncthreads();
  if (threadIdx.x < 32) {                                 // warp 0: exclusive prefix of the 256 counts, in place
    const int lane = threadIdx.x;
    int c[8], sum = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) { c[q] = s_hist[lane * 8 + q]; sum += c[q]; }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    int run = incl - sum;
#pragma unroll
    for (int q = 0; q < 8; ++q) { s_hist[lane * 8 + q] = run; run += c[q]; }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < count; k += NT) {         // scatter: the bucket's cursor ends at the bucket's end
    const int v = s_ids[k];
    s_tmp[atomicAdd(&s_hist[(v - mn) >> shift], 1)] = v;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < count; k += NT) {         // rank inside the bucket (ids are unique)
    const int x = s_tmp[k];
    const int b = (x - mn) >> shift;
    const int s = b > 0 ? s_hist[b - 1] : 0, e = s_hist[b];
    int r = 0;
    for (int j = s; j < e; ++j) r += (s_tmp[j] < x) ? 1 : 0;
    s_ids[s + r] = x;
  }
  __syncthreads();
}

// bits lo..hi (inclusive, 0 <= lo <= hi <= 15)
__device__ __forceinline__ unsigned bit_span16(int lo, int hi) { return ((2u << hi) - 1u) & ~((1u << lo) - 1u); }

// The work of one tile (tile_global = view * tiles_per_view + tile); called by the one-CTA-per-tile kernel and by
// the persistent overflow kernel below.
template <typename T, int MODE, int NT>
__device__ __forceinline__ void soft_tile(const SoftParams<T>& sp, const FaceLayout& fl, const Geometry& geo,
                                          const BinLists& L, const int b, const int tx, const int ty) {
  typedef typename CellInt<T>::type cell_int;
  typedef SoftSmem<T, NT> SM;
  constexpr int NW = NT / 32, CAP = SM::CAP, CELLS_TOT = SM::CELLS_TOT;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int* s_ids = reinterpret_cast<int*>(smem_raw + SM::ids);              // [CAP] sorted candidate faces
  unsigned* s_mask = reinterpret_cast<unsigned*>(smem_raw + SM::mask);  // [CAP] col bits | row bits << 16
  T* s_cell = reinterpret_cast<T*>(smem_raw + SM::cell);                // [CELLS_TOT] k-list cells
  int* s_upix = reinterpret_cast<int*>(smem_raw + SM::upix);            // [256] tile-local index of uncovered px
  int* s_wtab = reinterpret_cast<int*>(smem_raw + SM::wtab);            // [NW][32] per-warp owner tables
  int* s_ktot = reinterpret_cast<int*>(smem_raw + SM::ktot);            // [256] hits so far per uncovered px
  int* s_knew = reinterpret_cast<int*>(smem_raw + SM::knew);            // [256] new hits (this pass, super-batch)
  T* s_acc = reinterpret_cast<T*>(smem_raw + SM::acc);                  // [256] running product of (1 - prob)
  __shared__ int scratch[10];                                           // [0..7] warp sums, [8] append cursor, [9] U
  __shared__ float s_px[TILE], s_py[TILE];

  const unsigned tile_global = (unsigned)b * (unsigned)geo.tiles_per_view + (unsigned)(ty * geo.tiles_x + tx);
  const int rx0 = tx * TILE, ry0 = ty * TILE, rx1 = rx0 + TILE - 1, ry1 = ry0 + TILE - 1;
  const int64_t vbase = fl.base(b);
  const int K = sp.knum;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned lt_mask = (1u << lane) - 1u, le_mask = 0xffffffffu >> (31 - lane);

  // ---- phase A: covered pixels are final; the uncovered ones are compacted
  const bool from_map = MODE == MODE_FWD && sp.tile_unc != nullptr;    // kernel-uniform
  unsigned um[TILE_PIX / 32] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
  if (from_map) {
    // the rasterizer left the tile's coverage map and every pixel's provisional value: a tile without uncovered
    // pixels or without candidate faces is finished (most tiles: ~100 instructions instead of the face_idx sweep)
    const uint4* m4 = reinterpret_cast<const uint4*>(sp.tile_unc + (size_t)tile_global * (TILE_PIX / 32));
    const uint4 a = __ldg(m4), c = __ldg(m4 + 1);
    um[0] = a.x; um[1] = a.y; um[2] = a.z; um[3] = a.w; um[4] = c.x; um[5] = c.y; um[6] = c.z; um[7] = c.w;
    int tot = 0;
#pragma unroll
    for (int q = 0; q < TILE_PIX / 32; ++q) tot += __popc(um[q]);
    if (tot == 0) return;
    const TileLists t0 = tile_lists(L, geo, fl, b, tx, ty);
    if (t0.n_fine + t0.n_filtered() == 0) return;
  }
  if (threadIdx.x < TILE) {
    s_px[threadIdx.x] = pix_coord(pix_scale(sp.pix_mult, geo.W), geo.W, rx0 + threadIdx.x);
  } else if (threadIdx.x < 2 * TILE) {
    const int j = threadIdx.x - TILE;
    s_py[j] = pix_coord(pix_scale(sp.pix_mult, geo.H), geo.H, geo.H - 1 - (ry0 + j));
  }
  if (threadIdx.x == 0) scratch[9] = 0;
  __syncthreads();
  if (from_map) {
    for (int t = threadIdx.x; t < TILE_PIX; t += NT) {                 // warp w of the pass = word t >> 5 of the map
      unsigned mine = 0u; int before = 0;                              // (um[] is indexed with constants only)
#pragma unroll
      for (int q = 0; q < TILE_PIX / 32; ++q) {
        if (q == (t >> 5)) mine = um[q];
        if (q < (t >> 5)) before += __popc(um[q]);
      }
      if ((mine >> lane) & 1u) s_upix[before + __popc(mine & lt_mask)] = t;   // row-major order, no atomics
      s_ktot[t] = 0;
      s_acc[t] = (T)1.0;
    }
    if (threadIdx.x == 0) {
      int tot = 0;
#pragma unroll
      for (int q = 0; q < TILE_PIX / 32; ++q) tot += __popc(um[q]);
      scratch[9] = tot;
    }
  } else {
  for (int t = threadIdx.x; t < TILE_PIX; t += NT) {                   // uniform trip count
    const int px = rx0 + (t & (TILE - 1)), py = ry0 + (t / TILE);
    const bool in_image = px < geo.W && py < geo.H;
    const int64_t pix = ((int64_t)b * geo.H + py) * geo.W + px;
    const bool uncovered = in_image && (sp.face_idx[pix] < 0);
    if (MODE == MODE_FWD && in_image && !uncovered) {
      sp.soft[pix] = (T)1.0;                                            // :68-70
      if (sp.prob) {                                                    // dibr_soft_mask.cpp:86-96 defaults
        for (int k = 0; k < K; ++k) {
          sp.prob[pix * K + k] = (T)0;
          sp.close_idx[pix * K + k] = -1;
          sp.dist_type[pix * K + k] = 0;
        }
      }
    }
    const unsigned m = __ballot_sync(0xffffffffu, uncovered);
    if (m != 0u) {
      int base = 0;
      if (lane == 0) base = atomicAdd(&scratch[9], __popc(m));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (uncovered) s_upix[base + __popc(m & lt_mask)] = t;
    }
    s_ktot[t] = 0;                                       // hits so far of uncovered pixel u (< K)
    s_acc[t] = (T)1.0;                                   // forward: running product of (1 - prob), face order
  }
  }
  __syncthreads();
  const int U = scratch[9];
  if (U == 0) return;

  const TileLists tl = tile_lists(L, geo, fl, b, tx, ty);
  const int n_tile = tl.n_fine;
  const int* tile_list = tl.fine;
  const int n_filt = tl.n_filtered();
  const int nf = (int)(fl.base(b + 1) - vbase);
  const int KS = K | 1;                                  // odd row stride: conflict-free cell rows
  const int NPB = max(1, min(NT, CELLS_TOT / KS));       // pixels per super-batch (K > CELLS_TOT: 1, clipped rows)
  const int KROW = min(K, CELLS_TOT);                    // new hits a pixel can take per pass
  cell_int* s_celli = reinterpret_cast<cell_int*>(s_cell);
  int* w_tab = s_wtab + warp * 32;

  // The candidates are consumed in PASSES of increasing face-index ranges (lo, hi], each holding at most
  // SOFT_CAP candidates.  Nearly every tile needs a single pass (hi = last face); only very dense tiles
  // (config 5) take the selection path below.
  int lo = -1;
  bool first = true;
  // (with the distribution sort a pass costs little per candidate: the first pass takes as many as fit — measured
  //  at c3-64 with 256 / 512 / 1024 first: 1.644 / 1.585 / 1.580 ms; 128: 1.887)
  int target = sp.sel_first > 0 ? min(sp.sel_first, CAP) : CAP;
  while (true) {
    // ---- gather the candidates with lo < id <= hi: tile list + filtered large faces.  A pass takes about `target`
    //      of the lowest remaining ids (the k-lists are the FIRST knum hits in face order, so a pixel is usually
    //      complete long before the tile's whole list has been looked at); the target doubles from pass to pass.
    int hi = nf - 1;
    int count = 0;
    bool fits = first && n_tile + n_filt <= target;
    if (fits) {
      for (int k = threadIdx.x; k < n_tile; k += NT) s_ids[k] = tile_list[k];
      if (threadIdx.x == 0) scratch[8] = n_tile;
      __syncthreads();
      for (int base = 0; base < n_filt; base += NT) {                  // n_filt is CTA-uniform
        const int k = base + threadIdx.x;
        int f = -1;
        if (k < n_filt) {
          f = tl.filtered_at(k);
          if (!range_hits_rect(L.ranges[vbase + f], rx0, ry0, rx1, ry1)) f = -1;
        }
        const unsigned m = __ballot_sync(0xffffffffu, f >= 0);
        if (m != 0u) {                                                  // warp-aggregated append (sorted later)
          int pos = 0;
          if (lane == 0) pos = atomicAdd(&scratch[8], __popc(m));
          pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(m & lt_mask);
          if (f >= 0) s_ids[pos] = f;
        }
      }
      __syncthreads();
      count = scratch[8];
    } else {
      // -- selection: histogram over the remaining ids; a bisection on the face index only inside one bucket
      //    that alone exceeds the capacity (ids are unique, so a range of one id always fits)
      __syncthreads();
      int hi_bucket;
      hi = select_hi<NT>(tl, L.ranges + vbase, lo, nf, target, CAP, rx0, ry0, rx1, ry1,
                         reinterpret_cast<int*>(s_mask), scratch, count, hi_bucket);
      if (hi == lo) {
        int good = lo, bad = hi_bucket;
        while (bad - good > 1) {
          const int mid = good + ((bad - good) >> 1);
          if (count_candidates<NT>(tl, L.ranges + vbase, lo, mid, rx0, ry0, rx1, ry1, scratch) <= CAP) good = mid; else bad = mid;
        }
        hi = good;
      }
      if (threadIdx.x == 0) scratch[8] = 0;
      __syncthreads();
      for (int k = threadIdx.x; k < n_tile; k += NT) {
        const int f = tile_list[k];
        if (f > lo && f <= hi) s_ids[atomicAdd(&scratch[8], 1)] = f;
      }
      for (int k = threadIdx.x; k < n_filt; k += NT) {
        const int f = tl.filtered_at(k);
        if (f > lo && f <= hi && range_hits_rect(L.ranges[vbase + f], rx0, ry0, rx1, ry1))
          s_ids[atomicAdd(&scratch[8], 1)] = f;
      }
      __syncthreads();
      count = scratch[8];
    }
    target = min(2 * target, CAP);
    first = false;
    if (count > 0) {
      // ---- sort by face index; order is semantic (first knum faces in face order)
      __syncthreads();
      if (DIBR_SOFT_BUCKET_SORT > 0 && count > DIBR_SOFT_BUCKET_SORT) {
        // (s_mask is free until the masks are built below, s_knew between super-batches)
        bucket_sort<NT>(s_ids, reinterpret_cast<int*>(s_mask), s_knew, scratch, count);
      } else if (count <= 4 * NT) {
        chunk_rank_sort<NT>(s_ids, count);
      } else {
        int n2 = 1;
        while (n2 < count) n2 <<= 1;
        for (int k = count + threadIdx.x; k < n2; k += NT) s_ids[k] = 0x7fffffff;
        __syncthreads();
        for (int size = 2; size <= n2; size <<= 1) {
          for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = threadIdx.x; t < (n2 >> 1); t += NT) {
              const int lo_i = ((t / stride) * stride * 2) + (t % stride);
              const int hi_i = lo_i + stride;
              const bool up = ((lo_i & size) == 0);
              const int a = s_ids[lo_i], c = s_ids[hi_i];
              if ((a > c) == up) { s_ids[lo_i] = c; s_ids[hi_i] = a; }
            }
            __syncthreads();
          }
        }
      }
      // ---- tile-local coverage mask of every candidate: bit x (columns) and bit 16 + y (rows) of the pixels
      //      inside its enlarged bbox — the exact pixel range of the reference's float test (:89-97) — stored
      //      TRANSPOSED per chunk of 32 candidates: word 32 c + x = the candidates of chunk c that cover column x,
      //      word 32 c + 16 + y = those that cover row y, so that phase B tests 32 candidates against a pixel with
      //      one AND (a warp handles whole chunks: the loop index advances by NT, a multiple of 32)
      const int count32 = (count + 31) & ~31;
      for (int k = threadIdx.x; k < count32; k += NT) {
        unsigned m = 0u;
        if (k < count) {
          const uint2 r = L.ranges[vbase + s_ids[k]];
          const int lx0 = max((int)(r.x & 0xffffu), rx0) - rx0, lx1 = min((int)(r.x >> 16), rx1) - rx0;
          const int ly0 = max((int)(r.y & 0xffffu), ry0) - ry0, ly1 = min((int)(r.y >> 16), ry1) - ry0;
          if (lx0 <= lx1 && ly0 <= ly1) m = bit_span16(lx0, lx1) | (bit_span16(ly0, ly1) << 16);
        }
        s_mask[k] = warp_transpose32(m);
      }
      __syncthreads();

      // ---- phases B-D over super-batches of NPB uncovered pixels (usually the whole tile at once):
      //      B  lane = one of the warp's pixels: column word AND row word of every chunk of 32 candidates = the
      //         pixel's hits in that chunk; their slots are appended, in face order, to its cell row, until it
      //         has K in total
      //      C  each warp takes an equal slice of the super-batch's pixels; their new (pixel, k) pairs are
      //         flattened (prefix sum of the list lengths; owner by REDUX/ballot/popc) and evaluated one per
      //         lane with full lanes: distance, probability (forward: stored back into the cell; backward:
      //         gradients)
      //      D  forward, lane = one of the warp's pixels: multiplies the new probabilities, in face order, into the
      //         pixel's running product (:174-182)
      for (int sb0 = 0; sb0 < U; sb0 += NPB) {
        const int nsb = min(NPB, U - sb0);
        const int PW = (nsb + NW - 1) / NW;              // pixels per warp (<= 32) in phase C
        const int ubase = warp * PW;
        // -- B  (lane = one of the warp's own phase-C pixels; no CTA barrier before phase C)
        {
          const int npx = max(0, min(PW, nsb - ubase));
          const bool actb = lane < npx;
          const int u = ubase + lane;
          const int t = actb ? s_upix[sb0 + u] : 0;
          const int room = actb ? min(K - s_ktot[sb0 + u], KROW) : 0;
          cell_int* row = s_celli + u * KS;
          const unsigned* cm = s_mask + (t & 15);          // + 32 c: the chunk's candidates covering the pixel's column
          const unsigned* rm = s_mask + 16 + (t >> 4);     //         ... and its row
          int kid = 0;
          for (int c = 0; c < count32; c += 32) {          // candidates in face order: chunk by chunk, bit by bit
            if (!__any_sync(0xffffffffu, kid < room)) break;        // every list of the warp is full
            unsigned h = kid < room ? (cm[c] & rm[c]) : 0u;
            while (h != 0u && kid < room) {
              row[kid++] = (cell_int)(c + __ffs(h) - 1);
              h &= h - 1u;
            }
          }
          if (actb) s_knew[u] = kid;
        }
        __syncwarp();
        // -- C
        if (ubase < nsb) {
          const int npix = min(PW, nsb - ubase);
          const bool act = lane < npix;
          const int t = act ? s_upix[sb0 + ubase + lane] : 0;
          const int kid = act ? s_knew[ubase + lane] : 0;
          const int kprev = act ? s_ktot[sb0 + ubase + lane] : 0;
          int incl = kid;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
          }
          const int start = incl - kid;
          const int nq = __shfl_sync(0xffffffffu, incl, 31);
          const unsigned ne = __ballot_sync(0xffffffffu, kid > 0);
          if (kid > 0) w_tab[__popc(ne & lt_mask)] = lane;
          T g = 0, sft = 0;
          if (MODE == MODE_BWD && act) {
            const int64_t tpix = ((int64_t)b * geo.H + ry0 + (t >> 4)) * geo.W + rx0 + (t & 15);
            g = sp.grad_soft[tpix]; sft = sp.soft_in[tpix];
          }
          unsigned lbase = 0;
          const bool logging = MODE == MODE_FWD && sp.log_count != nullptr && nq > 0;
          const unsigned seg = (tile_global * 0x9E3779B1u) >> 26;     // hashed: neighbouring tiles (one silhouette)
                                                                      // spread over all LOG_SEGMENTS = 64 segments
          if (logging) {
            if (lane == 0) {
              lbase = atomicAdd(&sp.log_count[seg], (unsigned)nq);
              if (lbase + (unsigned)nq > sp.log_cap && atomicExch(&sp.tile_flag[tile_global], 1u) == 0u)
                sp.over_list[atomicAdd(sp.over_count, 1u)] = tile_global;      // first overflow of this tile
            }
            lbase = __shfl_sync(0xffffffffu, lbase, 0);
          }
          __syncwarp();
          const float prob_scale = __fdiv_rn(__fdiv_rn(sp.sigmainv, sp.pix_mult), sp.pix_mult);
          for (int q0 = 0; q0 < nq; q0 += 32) {
            const unsigned rel = (unsigned)(start - q0);
            const unsigned bit = (kid > 0 && rel < 32u) ? (1u << rel) : 0u;
            const unsigned starts = __reduce_or_sync(0xffffffffu, bit);
            const int before = __popc(__ballot_sync(0xffffffffu, kid > 0 && start < q0));
            const int rank = before + __popc(starts & le_mask) - 1;    // rank among the pixels with hits
            const int q = q0 + lane;
            const bool valid = q < nq;
            const int pl = valid ? w_tab[rank] : 0;
            const int pstart = __shfl_sync(0xffffffffu, start, pl);
            const int pt = __shfl_sync(0xffffffffu, t, pl);
            const int pk0 = __shfl_sync(0xffffffffu, kprev, pl);
            T pg = 0, ps = 0;
            if (MODE == MODE_BWD) { pg = __shfl_sync(0xffffffffu, g, pl); ps = __shfl_sync(0xffffffffu, sft, pl); }
            if (!valid) continue;
            const int k = q - pstart;
            const int cell = (ubase + pl) * KS + k;
            const int f = s_ids[(int)s_celli[cell]];
            const int lx = pt & 15, ly = pt >> 4;
            const T x0 = (T)s_px[lx], y0 = (T)s_py[ly];
            T p[6];
            load_face_xy<T>(sp.img, vbase + f, sp.coord_mult, sp.premultiplied, p);
            int edgeid;
            // fp32 without API-visible side buffers: reciprocal form of the per-edge divisions (dibr_common.cuh)
            // fp32 without API-visible side buffers: fp32 form of the per-edge divisions and of the exponent
            // (dibr_common.cuh: soft_distance_f32 / soft_prob_f32)
            T d, prob;
            soft_pair<T>(p, x0, y0, sp.pix_mult, sp.sigmainv, !sp.prob, prob_scale, d, prob, edgeid);
            if (MODE == MODE_FWD) {
              s_cell[cell] = prob;
              if (logging && lbase + (unsigned)q < sp.log_cap) {
                const size_t o = (size_t)seg * sp.log_cap + lbase + (unsigned)q;
                sp.log[o] = log_pack((unsigned)(((int64_t)b * geo.H + ry0 + ly) * geo.W + rx0 + lx),
                                     ((unsigned)(vbase + f) << 3) | (unsigned)edgeid, prob,
                                     (unsigned)(rx0 + lx) | ((unsigned)(ry0 + ly) << 16));
              }
              if (sp.prob) {
                const int64_t o = (((int64_t)b * geo.H + ry0 + ly) * geo.W + rx0 + lx) * K + pk0 + k;
                sp.prob[o] = prob; sp.close_idx[o] = (int64_t)f; sp.dist_type[o] = (uint8_t)(edgeid + 1);
              }
            } else {
              soft_backward_pair<T>(p, x0, y0, prob, edgeid, pg, ps, sp.sigmainv, sp.pix_mult,
                                    sp.grad_img + (vbase + f) * 6);
            }
          }
        }
        __syncwarp();
        // -- D  (lane = one of the warp's own phase-C pixels: the probabilities it reads were written by this warp)
        if (ubase + lane < nsb && lane < PW) {
          const int u = ubase + lane;
          const int kid = s_knew[u];
          if (MODE == MODE_FWD) {
            const T* row = s_cell + u * KS;
            T allprob = s_acc[sb0 + u];
            if (sizeof(T) == 4 && !sp.prob) {
              // fp32 fused path: the running product (:174-178, float * double rounded to float at every step in
              // the reference) is taken in fp32 — a dependent chain of fp64 conversions / multiplies per hit
              // otherwise; <= 1 ulp per factor, far inside the 1e-4 tolerance of the soft mask
              for (int k = 0; k < kid; ++k) allprob = allprob * ((T)1 - row[k]);
            } else {
              for (int k = 0; k < kid; ++k) allprob = (T)((double)allprob * (1.0 - (double)row[k]));
            }
            s_acc[sb0 + u] = allprob;
          }
          s_ktot[sb0 + u] += kid;
        }
        __syncthreads();                                 // cells / s_knew are reused by the next super-batch
      }
    }
    // ---- next pass?
    if (hi >= nf - 1) break;
    lo = hi;
    bool full = true;
    for (int u = threadIdx.x; u < U; u += NT) full = full && s_ktot[u] >= K;
    if (__syncthreads_and(full)) break;                                                // every list is full
  }
  __syncthreads();
  if (MODE == MODE_FWD) {
    for (int u = threadIdx.x; u < U; u += NT) {
      const int t = s_upix[u];
      const int64_t tpix = ((int64_t)b * geo.H + ry0 + (t >> 4)) * geo.W + rx0 + (t & 15);
      sp.soft[tpix] = (T)(1.0 - (double)s_acc[u]);
      if (sp.prob)
        for (int k = s_ktot[u]; k < K; ++k) {
          sp.prob[tpix * K + k] = (T)0; sp.close_idx[tpix * K + k] = -1; sp.dist_type[tpix * K + k] = 0;
        }
    }
  }
}

template <typename T, int MODE, int NT>
__global__ void __launch_bounds__(NT, (sizeof(T) == 4 ? DIBR_SOFT_TPS / NT : 512 / NT))
soft_mask_kernel(SoftParams<T> sp, FaceLayout fl, Geometry geo, BinLists L) {
  int b, tx, ty;
  if (!scheduled_tile(L, geo, b, tx, ty)) return;
  soft_tile<T, MODE, NT>(sp, fl, geo, L, b, tx, ty);
}

// Backward of the tiles whose pairs did not fit the log: a small persistent grid walks the overflow list
// (normally empty: the CTAs read one word and exit).
template <typename T, int NT>
__global__ void __launch_bounds__(NT, (sizeof(T) == 4 ? DIBR_SOFT_TPS / NT : 512 / NT))
soft_mask_overflow_kernel(SoftParams<T> sp, FaceLayout fl, Geometry geo, BinLists L) {
  const unsigned n = min(*sp.over_count, (unsigned)(geo.B * geo.tiles_per_view));
  for (unsigned i = blockIdx.x; i < n; i += gridDim.x) {
    const unsigned tg = sp.over_list[i];
    const int tile = (int)(tg % (unsigned)geo.tiles_per_view);
    soft_tile<T, MODE_BWD, NT>(sp, fl, geo, L, (int)(tg / (unsigned)geo.tiles_per_view), tile % geo.tiles_x,
                               tile / geo.tiles_x);
    __syncthreads();
  }
}

// ---------------------------------------------------------------- backward from the forward's pair log
// grid (LOG_CTAS_PER_SEGMENT, LOG_SEGMENTS); thread = logged (pixel, face) pair.  Same per-pair gradient as the
// recompute kernel, without re-deriving the k-lists: 12 bytes of log + one 24-byte face record per pair.
// One logged pair: everything that must be fetched before the arithmetic (issued for two pairs at a time so that
// the three dependent gathers — record -> pixel values / face record — overlap; ncu showed the kernel waiting on
// long-scoreboard, not on the atomics).
template <typename T>
struct LogItem {
  uint4 rec; T g, soft; unsigned flag, b, px, py; T p[6]; int64_t gf;
};
template <typename T>
__device__ __forceinline__ void log_pixel_coords(const uint4& rec, const Geometry& geo, const unsigned hw, unsigned& b,
                                                 unsigned& px, unsigned& py) {
  b = rec.x / hw;
  const unsigned rem = rec.x - b * hw;
  py = rem / (unsigned)geo.W; px = rem - py * (unsigned)geo.W;
}
template <>
__device__ __forceinline__ void log_pixel_coords<float>(const uint4& rec, const Geometry&, const unsigned hw,
                                                        unsigned& b, unsigned& px, unsigned& py) {
  b = 0u;                                   // only needed for the overflow flags: taken there
  px = rec.w & 0xffffu; py = rec.w >> 16;
}
// `any_over`: some tile overflowed the log (normally none: the per-pair flag gather is skipped)
template <typename T>
__device__ __forceinline__ void log_fetch_pixel(const SoftParams<T>& sp, const Geometry& geo, const unsigned hw,
                                                const bool any_over, LogItem<T>& it) {
  const unsigned pix = it.rec.x;
  log_pixel_coords<T>(it.rec, geo, hw, it.b, it.px, it.py);
  it.g = sp.grad_soft[pix];
  it.soft = sp.soft_in[pix];
  it.flag = 0u;
  if (any_over) {                           // rare: some tile overflowed the log
    if (sizeof(T) == 4) it.b = pix / hw;
    it.flag = sp.tile_flag[it.b * geo.tiles_per_view + (it.py / TILE) * geo.tiles_x + it.px / TILE];
  }
}
template <typename T>
__device__ __forceinline__ void log_fetch_face(const SoftParams<T>& sp, const FaceLayout& fl, LogItem<T>& it) {
  it.gf = (int64_t)(it.rec.y >> 3);         // records carry the face's index in the whole batch (view * F + face)
  load_face_xy<T>(sp.img, it.gf, sp.coord_mult, sp.premultiplied, it.p);
}

#ifndef DIBR_LOG_INFLIGHT
#define DIBR_LOG_INFLIGHT 2       // logged pairs whose gathers are in flight together, per thread
#endif
#ifndef DIBR_LOG_CTAS
#define DIBR_LOG_CTAS 6           // resident 256-thread CTAs per SM the log backward's register allocation aims at
#endif
template <typename T>
__global__ void __launch_bounds__(256, (sizeof(T) == 4 ? DIBR_LOG_CTAS : 2)) soft_mask_bwd_log_kernel(SoftParams<T> sp, FaceLayout fl, Geometry geo) {
  const unsigned seg = blockIdx.y;
  const unsigned n = min(sp.log_count[seg], sp.log_cap);
  const unsigned hw = (unsigned)geo.H * (unsigned)geo.W;
  const bool any_over = *sp.over_count != 0u;
  const float sx = pix_scale(sp.pix_mult, geo.W), sy = pix_scale(sp.pix_mult, geo.H);
  const uint4* log = sp.log + (size_t)seg * sp.log_cap;
  const unsigned stride = gridDim.x * blockDim.x;
  // (summing lanes that hold the same face with shuffles before the atomics was measured slower: c2 0.121 vs
  // 0.108 ms, c3 0.175 vs 0.154 ms — the kernel waits on its gathers, not on the atomics)
  for (unsigned i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += DIBR_LOG_INFLIGHT * stride) {
    LogItem<T> it[DIBR_LOG_INFLIGHT];
    bool on[DIBR_LOG_INFLIGHT];
#pragma unroll
    for (int q = 0; q < DIBR_LOG_INFLIGHT; ++q) {
      const unsigned j = i0 + (unsigned)q * stride;
      on[q] = j < n;
      it[q].rec = log[on[q] ? j : i0];
    }
#pragma unroll
    for (int q = 0; q < DIBR_LOG_INFLIGHT; ++q) log_fetch_pixel<T>(sp, geo, hw, any_over, it[q]);
#pragma unroll
    for (int q = 0; q < DIBR_LOG_INFLIGHT; ++q) log_fetch_face<T>(sp, fl, it[q]);
#pragma unroll
    for (int q = 0; q < DIBR_LOG_INFLIGHT; ++q) {
      if (on[q] && it[q].g != (T)0 && !it[q].flag) {
        const T x0 = (T)pix_coord(sx, geo.W, (int)it[q].px), y0 = (T)pix_coord(sy, geo.H, geo.H - 1 - (int)it[q].py);
        soft_backward_pair<T>(it[q].p, x0, y0, log_prob<T>(it[q].rec), (int)(it[q].rec.y & 7u), it[q].g, it[q].soft,
                              sp.sigmainv, sp.pix_mult, sp.grad_img + it[q].gf * 6);
      }
    }
  }
}

// ---------------------------------------------------------------- reference-shaped backward from side buffers
template <typename T>
__global__ void __launch_bounds__(256) soft_mask_bwd_sidebuf_kernel(
    const T* __restrict__ grad_soft, const T* __restrict__ soft, const int64_t* __restrict__ face_idx,
    const T* __restrict__ prob, const int64_t* __restrict__ close_idx, const uint8_t* __restrict__ dist_type,
    const T* __restrict__ img_mult, T* grad_img, int B, int H, int W, int F, int K, float sigmainv,
    float multiplier) {
  const int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= (int64_t)B * H * W) return;
  if (face_idx[pix] >= 0) return;                                       // :266
  const int px = (int)(pix % W);
  const int py = (int)((pix / W) % H);
  const int b = (int)(pix / ((int64_t)W * H));
  const T x0 = (T)pix_coord(pix_scale(multiplier, W), W, px);
  const T y0 = (T)pix_coord(pix_scale(multiplier, H), H, H - 1 - py);
  const T dLdp = grad_soft[pix], allprob = soft[pix];
  for (int k = 0; k < K; ++k) {                                         // :271-276
    const int64_t f = close_idx[pix * K + k];
    if (f < 0) break;
    if (f >= F) continue;
    const int64_t s6 = ((int64_t)b * F + f) * 6;
    T p[6];
#pragma unroll
    for (int q = 0; q < 6; ++q) p[q] = img_mult[s6 + q];
    soft_backward_pair<T>(p, x0, y0, prob[pix * K + k], (int)dist_type[pix * K + k] - 1, dLdp, allprob,
                          sigmainv, multiplier, grad_img + s6);
  }
}

}  // namespace dibr
