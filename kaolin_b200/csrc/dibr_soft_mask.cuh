// dibr_soft_mask.cuh — K3/K5: DIB-R soft mask forward and backward (reference:
// dibr_soft_mask_cuda.cu:27-184 and :230-353).
//
// One CTA = one 16x16 tile.  Only tiles with at least one uncovered pixel do any work beyond the
// face_idx read / soft write.  The tile's candidate faces (enlarged-bbox tile list + filtered "large"
// faces) are gathered into shared memory, bitonic-sorted by face index — the reference keeps the FIRST
// `knum` faces in face order whose enlarged bbox contains the pixel (:80-171), so order is semantic —
// and their pixel ranges are staged next to them.  Each uncovered pixel then walks the sorted list.
//
// The walk is split in phases so that the expensive part runs with full lanes even when only a few
// pixels of the tile are uncovered: (A) compact the uncovered pixels; then, warp-locally, each uncovered pixel
// of a warp's batch (B) gets its k-list by ballot/popc over 32 candidates at a time, appended to the warp's
// pair queue, (C) the queued (pixel, k) pairs are evaluated one per lane, (D) one lane per pixel
// multiplies the probabilities in face order.
//
// MODE_FWD writes soft (+ optional k-list side buffers, API-visible through
// _C.render.mesh.dibr_soft_mask_forward_cuda).  MODE_BWD re-derives the same k-list with the same
// arithmetic instead of reading 13*knum bytes per pixel of side buffers, and accumulates the vertex
// gradients with red.global.add.
#pragma once
#include "dibr_binning.cuh"

namespace dibr {

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
  // forward
  T* soft;                    // (B,H,W)
  T* prob;                    // (B,H,W,K) or nullptr
  int64_t* close_idx;         // (B,H,W,K) or nullptr
  uint8_t* dist_type;         // (B,H,W,K) or nullptr
  // backward
  const T* grad_soft;         // (B,H,W)
  const T* soft_in;           // (B,H,W)
  T* grad_img;                // (B,F,3,2)
};

// dibr_soft_mask_cuda.cu:283-347 for one (pixel, face) pair; p = multiplied coordinates.
template <typename T>
__device__ __forceinline__ void soft_backward_pair(const T* p, const T x0, const T y0, const T prob,
                                                   const int edgeid, const T dLdp, const T allprob,
                                                   const float sigmainv, const float multiplier, T* gi) {
  const double EPS = 1e-7;
  const T dLdz = (T)(-1.0 * (double)sigmainv * (double)dLdp * (1.0 - (double)allprob) /
                     (1.0 - (double)prob + EPS) * (double)prob);
  if (edgeid >= 3) {
    const int o = (edgeid - 3) * 2;
    const T x1 = p[o], y1 = p[o + 1];
    atomicAdd(gi + o, dLdz * 2 * (x1 - x0) / (T)multiplier);
    atomicAdd(gi + o + 1, dLdz * 2 * (y1 - y0) / (T)multiplier);
  } else {
    const int o = edgeid * 2, o2 = ((edgeid + 1) % 3) * 2;
    const T x1 = p[o], y1 = p[o + 1], x2 = p[o2], y2 = p[o2 + 1];
    const T A = y2 - y1, Bc = x1 - x2, Cc = x2 * y1 - x1 * y2;
    const T up = A * x0 + Bc * y0 + Cc;
    const T down = A * A + Bc * Bc;
    const double den = (double)down + EPS;
    const T dissquare = (T)((double)(up * up) / den);
    const T dzdA = (T)((double)(2 * (x0 * up - dissquare * A)) / den);
    const T dzdB = (T)((double)(2 * (y0 * up - dissquare * Bc)) / den);
    const T dzdC = (T)((double)(2 * up) / den);
    atomicAdd(gi + o, dLdz * (dzdB - y2 * dzdC) / (T)multiplier);
    atomicAdd(gi + o + 1, dLdz * (x2 * dzdC - dzdA) / (T)multiplier);
    atomicAdd(gi + o2, dLdz * (y1 * dzdC - dzdB) / (T)multiplier);
    atomicAdd(gi + o2 + 1, dLdz * (dzdA - x1 * dzdC) / (T)multiplier);
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

// Shared-memory plan (dynamic): candidates (ids + ranges), the compacted uncovered pixels, and the
// per-pixel k-lists of the current pixel chunk (PAIR_CAP (pixel, k) slots).
constexpr int PAIR_CAP = 2048;     // k-list slots: 8 warps x WARP_K
constexpr int WARP_K = PAIR_CAP / 8; // largest knum served by the fast path
template <typename T>
struct SoftSmem {
  static constexpr size_t ids = 0;
  static constexpr size_t rng = ids + (SOFT_CAP + 8) * sizeof(int);   // + INT_MAX padding
  static constexpr size_t hit = rng + SOFT_CAP * sizeof(uint2);
  static constexpr size_t prob = hit + PAIR_CAP * sizeof(int);
  static constexpr size_t upix = prob + PAIR_CAP * sizeof(T);
  static constexpr size_t pt = upix + 256 * sizeof(int);
  static constexpr size_t start = pt + 8 * 32 * sizeof(int);
  static constexpr size_t own = start + 8 * 33 * sizeof(int);
  static constexpr size_t total = own + PAIR_CAP;
};

// Each thread ranks Q of the ids (k = tid + q*256); see the call site.
template <int Q>
__device__ __forceinline__ void rank_sort(int* s_ids, const int count) {
  int mine[Q], rank[Q];
  bool any = false;
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const int k = threadIdx.x + q * 256;
    mine[q] = (k < count) ? s_ids[k] : 0x7fffffff;
    rank[q] = 0;
    any |= k < count;
  }
  if (any) {
    const int4* v4 = reinterpret_cast<const int4*>(s_ids);
    const int n4 = (count + 3) >> 2;
    for (int j = 0; j < n4; ++j) {
      const int4 v = v4[j];                                              // broadcast read
#pragma unroll
      for (int q = 0; q < Q; ++q)
        rank[q] += (v.x < mine[q]) + (v.y < mine[q]) + (v.z < mine[q]) + (v.w < mine[q]);
    }
  }
  __syncthreads();
#pragma unroll
  for (int q = 0; q < Q; ++q)
    if (threadIdx.x + q * 256 < count) s_ids[rank[q]] = mine[q];
  __syncthreads();
}

template <typename T, int MODE>
__global__ void __launch_bounds__(256) soft_mask_kernel(SoftParams<T> sp, FaceLayout fl, Geometry geo,
                                                        BinLists L) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int* s_ids = reinterpret_cast<int*>(smem_raw + SoftSmem<T>::ids);      // [SOFT_CAP] sorted candidate faces
  uint2* s_rng = reinterpret_cast<uint2*>(smem_raw + SoftSmem<T>::rng);  // [SOFT_CAP] their pixel ranges
  int* s_hit = reinterpret_cast<int*>(smem_raw + SoftSmem<T>::hit);      // [PAIR_CAP] k-lists, [u*K + k]
  T* s_prob = reinterpret_cast<T*>(smem_raw + SoftSmem<T>::prob);        // [PAIR_CAP] forward probabilities
  int* s_upix = reinterpret_cast<int*>(smem_raw + SoftSmem<T>::upix);    // [256] thread index of uncovered px
  int* s_pt = reinterpret_cast<int*>(smem_raw + SoftSmem<T>::pt);
  int* s_start = reinterpret_cast<int*>(smem_raw + SoftSmem<T>::start);
  unsigned char* s_own = smem_raw + SoftSmem<T>::own;
  __shared__ int scratch[9];

  const int tile = blockIdx.x % geo.tiles_per_view;
  const int b = blockIdx.x / geo.tiles_per_view;
  const int tx = tile % geo.tiles_x, ty = tile / geo.tiles_x;
  const int px = tx * TILE + (threadIdx.x & (TILE - 1));
  const int py = ty * TILE + (threadIdx.x / TILE);
  const bool in_image = px < geo.W && py < geo.H;
  const int64_t vbase = fl.base(b);
  const int64_t pix = ((int64_t)b * geo.H + py) * geo.W + px;
  const bool uncovered = in_image && (sp.face_idx[pix] < 0);
  const int K = sp.knum;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  if (MODE == MODE_FWD && in_image && !uncovered) {
    sp.soft[pix] = (T)1.0;                                              // :68-70
    if (sp.prob) {                                                      // dibr_soft_mask.cpp:86-96 defaults
      for (int k = 0; k < K; ++k) {
        sp.prob[pix * K + k] = (T)0;
        sp.close_idx[pix * K + k] = -1;
        sp.dist_type[pix * K + k] = 0;
      }
    }
  }
  // ---- phase A: compact the uncovered pixels of the tile
  int U;
  const int uslot = block_compact_256(uncovered, scratch, U);
  if (U == 0) return;
  if (uncovered) s_upix[uslot] = threadIdx.x;

  // ---- gather the candidates: tile list + filtered large faces
  const int n_tile = L.tile_count[blockIdx.x];
  const int* tile_list = L.pairs + L.tile_offset[blockIdx.x];
  const int n_large = L.large_count[b];
  const int* large_list = L.large_list + vbase;
  const int rx0 = tx * TILE, ry0 = ty * TILE, rx1 = rx0 + TILE - 1, ry1 = ry0 + TILE - 1;
  bool overflow = n_tile > SOFT_CAP || K > WARP_K;
  int count = 0;
  if (!overflow) {
    for (int k = threadIdx.x; k < n_tile; k += 256) s_ids[k] = tile_list[k];
    count = n_tile;
    for (int base = 0; base < n_large; base += 256) {                  // n_large is CTA-uniform
      const int k = base + threadIdx.x;
      int f = -1;
      if (k < n_large) {
        f = large_list[k];
        if (!range_hits_rect(L.ranges[vbase + f], rx0, ry0, rx1, ry1)) f = -1;
      }
      int add;
      const int slot = block_compact_256(f >= 0, scratch, add);
      if (count + add > SOFT_CAP) { overflow = true; break; }          // uniform
      if (f >= 0) s_ids[count + slot] = f;
      count += add;
    }
  }

  if (overflow) {
    // ---- rare: more candidates than shared memory holds -> the reference's in-order scan over every face
    if (uncovered) {
      SoftPixel<T, MODE> st;
      st.pix = pix; st.kid = 0; st.allprob = (T)1.0;
      st.x0 = (T)pix_coord(pix_scale(sp.pix_mult, geo.W), geo.W, px);
      st.y0 = (T)pix_coord(pix_scale(sp.pix_mult, geo.H), geo.H, geo.H - 1 - py);
      if (MODE == MODE_BWD) { st.dLdp = sp.grad_soft[pix]; st.soft = sp.soft_in[pix]; }
      const int nf = (int)(fl.base(b + 1) - vbase);
      for (int f = 0; f < nf && st.kid < K; ++f) {
        if (!range_hits_pixel(L.ranges[vbase + f], px, py)) continue;
        soft_visit<T, MODE>(sp, st, vbase, f);
      }
      if (MODE == MODE_FWD) {
        sp.soft[pix] = (T)(1.0 - (double)st.allprob);
        if (sp.prob)
          for (int k = st.kid; k < K; ++k) {
            sp.prob[pix * K + k] = (T)0; sp.close_idx[pix * K + k] = -1; sp.dist_type[pix * K + k] = 0;
          }
      }
    }
    return;
  }

  // ---- sort by face index; order is semantic (first knum faces in face order)
  __syncthreads();
  if (count <= 1024) {
    // rank sort: face ids are unique, so rank = number of smaller ids; one pass, one barrier.  The list is
    // padded to a multiple of 4 with INT_MAX (never smaller than anything) and read as int4 broadcasts.
    if (threadIdx.x < 4) s_ids[count + threadIdx.x] = 0x7fffffff;       // SOFT_CAP + 4 slots are reserved
    __syncthreads();
    if (count <= 256) rank_sort<1>(s_ids, count);
    else if (count <= 512) rank_sort<2>(s_ids, count);
    else rank_sort<4>(s_ids, count);
  } else {
    int n2 = 1;
    while (n2 < count) n2 <<= 1;
    for (int k = count + threadIdx.x; k < n2; k += 256) s_ids[k] = 0x7fffffff;
    __syncthreads();
    for (int size = 2; size <= n2; size <<= 1) {
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
        for (int t = threadIdx.x; t < (n2 >> 1); t += 256) {
          const int lo = ((t / stride) * stride * 2) + (t % stride);
          const int hi = lo + stride;
          const bool up = ((lo & size) == 0);
          const int a = s_ids[lo], c = s_ids[hi];
          if ((a > c) == up) { s_ids[lo] = c; s_ids[hi] = a; }
        }
        __syncthreads();
      }
    }
  }
  for (int k = threadIdx.x; k < count; k += 256) s_rng[k] = L.ranges[vbase + s_ids[k]];
  __syncthreads();

  // ---- phases B-D are warp-local (no block barriers).  Each warp owns the uncovered pixels u = warp,
  //      warp+8, ... and works in batches of up to 32 pixels / WARP_K (pixel, k) pairs:
  //      B  per pixel, ballot/popc over 32 candidates at a time -> its k-list (first K hits in face
  //         order) appended to the warp's pair queue
  //      C  the queued pairs are evaluated one per lane with full lanes (distance, probability;
  //         backward: gradients) — pixels with few hits share a warp pass
  //      D  one lane per pixel multiplies its probabilities in face order (:174-182)
  const float sx = pix_scale(sp.pix_mult, geo.W), sy = pix_scale(sp.pix_mult, geo.H);
  int* w_hit = s_hit + warp * WARP_K;                 // queued face ids
  T* w_prob = s_prob + warp * WARP_K;                 // their probabilities (forward)
  unsigned char* w_own = s_own + warp * WARP_K;       // batch-local pixel slot of each queued pair
  int* w_pt = s_pt + warp * 32;                       // per batch pixel: thread index in the tile
  int* w_start = s_start + warp * 33;                 // per batch pixel: first queue position (+ end)
  int u = warp;
  while (u < U) {
    int npix = 0, nq = 0;
    // -- B: fill the batch
    while (u < U && npix < 32) {
      const int t = s_upix[u];
      const int qx = tx * TILE + (t & (TILE - 1)), qy = ty * TILE + (t / TILE);
      // stop if this pixel's list might not fit (its length is at most min(K, count))
      if (nq + min(K, count) > WARP_K) break;
      int kid = 0;
      for (int base = 0; base < count && kid < K; base += 32) {
        const int c = base + lane;
        const bool hit = c < count && range_hits_pixel(s_rng[c], qx, qy);
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        const int r = kid + __popc(m & ((1u << lane) - 1u));
        if (hit && r < K) { w_hit[nq + r] = s_ids[c]; w_own[nq + r] = (unsigned char)npix; }
        kid = min(K, kid + __popc(m));
      }
      if (lane == 0) { w_pt[npix] = t; w_start[npix] = nq; }
      nq += kid;
      ++npix;
      u += 8;
    }
    if (lane == 0) w_start[npix] = nq;
    __syncwarp();
    // -- C: all queued pairs, one per lane
    for (int q = lane; q < nq; q += 32) {
      const int slot = w_own[q];
      const int t = w_pt[slot];
      const int qx = tx * TILE + (t & (TILE - 1)), qy = ty * TILE + (t / TILE);
      const T x0 = (T)pix_coord(sx, geo.W, qx);
      const T y0 = (T)pix_coord(sy, geo.H, geo.H - 1 - qy);
      const int f = w_hit[q];
      T p[6];
      load_face_xy<T>(sp.img, vbase + f, sp.coord_mult, sp.premultiplied, p);
      int edgeid;
      const T d = soft_distance<T>(p, x0, y0, sp.pix_mult, edgeid);
      const T prob = soft_prob<T>(d, sp.sigmainv, sp.pix_mult);
      const int64_t qpix = ((int64_t)b * geo.H + qy) * geo.W + qx;
      if (MODE == MODE_FWD) {
        w_prob[q] = prob;
        if (sp.prob) {
          const int64_t o = qpix * K + (q - w_start[slot]);
          sp.prob[o] = prob; sp.close_idx[o] = (int64_t)f; sp.dist_type[o] = (uint8_t)(edgeid + 1);
        }
      } else {
        soft_backward_pair<T>(p, x0, y0, prob, edgeid, sp.grad_soft[qpix], sp.soft_in[qpix], sp.sigmainv,
                              sp.pix_mult, sp.grad_img + (vbase + f) * 6);
      }
    }
    if (MODE == MODE_FWD) {
      __syncwarp();
      // -- D: one lane per pixel of the batch
      if (lane < npix) {
        const int t = w_pt[lane];
        const int qx = tx * TILE + (t & (TILE - 1)), qy = ty * TILE + (t / TILE);
        const int64_t qpix = ((int64_t)b * geo.H + qy) * geo.W + qx;
        const int q0 = w_start[lane], q1 = w_start[lane + 1];
        T allprob = (T)1.0;
        for (int q = q0; q < q1; ++q) allprob = (T)((double)allprob * (1.0 - (double)w_prob[q]));
        sp.soft[qpix] = (T)(1.0 - (double)allprob);
        if (sp.prob)
          for (int k = q1 - q0; k < K; ++k) {
            sp.prob[qpix * K + k] = (T)0; sp.close_idx[qpix * K + k] = -1; sp.dist_type[qpix * K + k] = 0;
          }
      }
    }
    __syncwarp();
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
