// dibr_texture.cuh — the step immediately AFTER the rasterization path:
// texture_mapping of kaolin/render/mesh/utils.py:23-76 — clamp uv to [0,1], map to [-1,1], flip y,
// torch.nn.functional.grid_sample(mode = nearest | bilinear, align_corners = False, padding_mode = 'border'),
// permute to channels-last — optionally followed by the tutorial's epilogue
// `torch.clamp(image * mask, 0., 1.)` (examples/tutorial/ian_dibr.py:254-258), in ONE kernel per direction:
// thread = output point; reads uv (8 B) [+ mask], gathers the 4 (or 1) texels per channel from the planar
// (C,Ht,Wt) texture and writes the C channels of the point contiguously (no (B,C,N,1) intermediate, no permute,
// no separate mul / clamp passes).  The texture may be shared by the whole batch (Bt == 1), which removes the
// `texture_map.repeat(batch_size, 1, 1, 1)` copy of the tutorial.
// The sampling arithmetic follows ATen's grid sampler (unnormalise ((g+1)*size-1)/2, clip to [0,size-1],
// nw/ne/sw/se weights from the floor corner, nearbyint for nearest) so that the reference's known-answer tests
// (tests/python/kaolin/render/mesh/test_utils.py:61-112) hold with torch.equal; in general parity is 1e-5 relative.
#pragma once
#include "dibr_common.cuh"

namespace dibr {

enum { TEX_NEAREST = 0, TEX_BILINEAR = 1 };

template <typename T>
struct TextureParams {
  const T* uv;        // (B,N,2) in [0,1] OpenGL convention (y up)
  const T* tex;       // (Bt,C,Ht,Wt), Bt = 1 or B
  const T* mask;      // (B,N) or nullptr: out = clamp(sample * mask, 0, 1) when given
  int B, Bt, C, Ht, Wt;
  int64_t N;
  int mode;
  T* out;             // (B,N,C)
  // backward
  const T* g_out;     // (B,N,C)
  T* g_uv;            // (B,N,2) or nullptr
  T* g_tex;           // (Bt,C,Ht,Wt) or nullptr, zero-filled by the host wrapper
  T* g_mask;          // (B,N) or nullptr
};

template <typename T>
struct TexTaps {
  int x0, y0, x1, y1;          // nw corner and se corner (clipped copies used for addressing)
  bool vx0, vx1, vy0, vy1;     // within bounds
  T wnw, wne, wsw, wse;
  T ix, iy;                    // clipped source coordinates
  T dix, diy;                  // d(ix)/d(u), d(iy)/d(v) including clamp, flip, unnormalise and clip
};

template <typename T> __device__ __forceinline__ T floor_t(T v);
template <> __device__ __forceinline__ float floor_t<float>(float v) { return floorf(v); }
template <> __device__ __forceinline__ double floor_t<double>(double v) { return floor(v); }
template <typename T> __device__ __forceinline__ T nearby_t(T v);
template <> __device__ __forceinline__ float nearby_t<float>(float v) { return nearbyintf(v); }
template <> __device__ __forceinline__ double nearby_t<double>(double v) { return nearbyint(v); }

// uv component -> clipped source index (ATen GridSamplerKernel: grid_sampler_compute_source_index with
// align_corners = false, padding border) and its derivative w.r.t. the ORIGINAL [0,1] coordinate.
template <typename T>
__device__ __forceinline__ T source_index(T c, const int size, const bool flip, T& deriv) {
  T d = (T)1;
  if (c < (T)0) { c = (T)0; d = (T)0; }              // torch.clamp(.., 0., 1.), utils.py:62 (NaN passes through)
  if (c > (T)1) { c = (T)1; d = (T)0; }
  T g = c * (T)2 - (T)1;                             // utils.py:63
  d *= (T)2;
  if (flip) { g = -g; d = -d; }                      // utils.py:64
  T x = ((g + (T)1) * (T)size - (T)1) / (T)2;        // unnormalise, align_corners = false
  d *= (T)size / (T)2;
  const T hi = (T)(size - 1);                        // clip_coordinates_set_grad
  if (!(x > (T)0)) { x = (T)0; d = (T)0; }           // also NaN -> 0, as ::max(in, 0) does in the forward of ATen
  else if (x >= hi) { x = hi; d = (T)0; }
  deriv = d;
  return x;
}

template <typename T>
__device__ __forceinline__ TexTaps<T> make_taps(const T u, const T v, const int Wt, const int Ht) {
  TexTaps<T> t;
  t.ix = source_index<T>(u, Wt, false, t.dix);
  t.iy = source_index<T>(v, Ht, true, t.diy);
  const T fx = floor_t<T>(t.ix), fy = floor_t<T>(t.iy);
  t.x0 = (int)fx; t.y0 = (int)fy; t.x1 = t.x0 + 1; t.y1 = t.y0 + 1;
  const T x1f = fx + (T)1, y1f = fy + (T)1;
  t.wnw = (x1f - t.ix) * (y1f - t.iy);
  t.wne = (t.ix - fx) * (y1f - t.iy);
  t.wsw = (x1f - t.ix) * (t.iy - fy);
  t.wse = (t.ix - fx) * (t.iy - fy);
  t.vx0 = t.x0 >= 0 && t.x0 < Wt; t.vx1 = t.x1 >= 0 && t.x1 < Wt;
  t.vy0 = t.y0 >= 0 && t.y0 < Ht; t.vy1 = t.y1 >= 0 && t.y1 < Ht;
  return t;
}

template <typename T>
__global__ void __launch_bounds__(256) texture_mapping_fwd_kernel(TextureParams<T> tp) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (int64_t)tp.B * tp.N) return;
  const int b = (int)(g / tp.N);
  const T u = tp.uv[g * 2], v = tp.uv[g * 2 + 1];
  const int64_t plane = (int64_t)tp.Ht * tp.Wt;
  const T* tex = tp.tex + (tp.Bt == 1 ? (int64_t)0 : (int64_t)b * tp.C * plane);
  const bool shade = tp.mask != nullptr;
  const T m = shade ? tp.mask[g] : (T)1;
  T* out = tp.out + g * tp.C;
  if (tp.mode == TEX_NEAREST) {
    T dx, dy;
    const T ix = source_index<T>(u, tp.Wt, false, dx), iy = source_index<T>(v, tp.Ht, true, dy);
    const int x = (int)nearby_t<T>(ix), y = (int)nearby_t<T>(iy);
    const bool ok = x >= 0 && x < tp.Wt && y >= 0 && y < tp.Ht;      // false for NaN coordinates
    for (int c = 0; c < tp.C; ++c) {
      T val = ok ? tex[c * plane + (int64_t)y * tp.Wt + x] : (T)0;
      if (shade) { val *= m; val = val < (T)0 ? (T)0 : (val > (T)1 ? (T)1 : val); }
      out[c] = val;
    }
    return;
  }
  const TexTaps<T> t = make_taps<T>(u, v, tp.Wt, tp.Ht);
  const int64_t o_nw = (int64_t)t.y0 * tp.Wt + t.x0, o_sw = (int64_t)t.y1 * tp.Wt + t.x0;
  for (int c = 0; c < tp.C; ++c) {
    const T* p = tex + c * plane;
    T acc = (T)0;
    if (t.vx0 && t.vy0) acc += p[o_nw] * t.wnw;
    if (t.vx1 && t.vy0) acc += p[o_nw + 1] * t.wne;
    if (t.vx0 && t.vy1) acc += p[o_sw] * t.wsw;
    if (t.vx1 && t.vy1) acc += p[o_sw + 1] * t.wse;
    if (shade) { acc *= m; acc = acc < (T)0 ? (T)0 : (acc > (T)1 ? (T)1 : acc); }
    out[c] = acc;
  }
}

// Backward.  g_tex gets red.global.add per tap and channel (neighbouring pixels of a rendered surface hit
// neighbouring texels: contention is low); g_uv is the bilinear derivative chained through clip / unnormalise /
// flip / clamp (zero for nearest, as in torch); with the shading epilogue the upstream gradient is first passed
// through clamp (inclusive bounds, torch.clamp backward) and the mask product.
template <typename T>
__global__ void __launch_bounds__(256) texture_mapping_bwd_kernel(TextureParams<T> tp) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (int64_t)tp.B * tp.N) return;
  const int b = (int)(g / tp.N);
  const T u = tp.uv[g * 2], v = tp.uv[g * 2 + 1];
  const int64_t plane = (int64_t)tp.Ht * tp.Wt;
  const int64_t toff = tp.Bt == 1 ? (int64_t)0 : (int64_t)b * tp.C * plane;
  const T* tex = tp.tex + toff;
  T* gtex = tp.g_tex ? tp.g_tex + toff : nullptr;
  const bool shade = tp.mask != nullptr;
  const T m = shade ? tp.mask[g] : (T)1;
  const T* go = tp.g_out + g * tp.C;
  T gm = (T)0;
  if (tp.mode == TEX_NEAREST) {
    T dx, dy;
    const T ix = source_index<T>(u, tp.Wt, false, dx), iy = source_index<T>(v, tp.Ht, true, dy);
    const int x = (int)nearby_t<T>(ix), y = (int)nearby_t<T>(iy);
    const bool ok = x >= 0 && x < tp.Wt && y >= 0 && y < tp.Ht;
    for (int c = 0; c < tp.C; ++c) {
      T gv = go[c];
      const int64_t o = c * plane + (int64_t)y * tp.Wt + x;
      if (shade) {
        const T s = ok ? tex[o] : (T)0;
        const T pre = s * m;
        if (!(pre >= (T)0 && pre <= (T)1)) gv = (T)0;
        gm += gv * s;
        gv *= m;
      }
      if (gtex && ok && gv != (T)0) atomicAdd(gtex + o, gv);
    }
    if (tp.g_uv) { tp.g_uv[g * 2] = (T)0; tp.g_uv[g * 2 + 1] = (T)0; }
    if (tp.g_mask) tp.g_mask[g] = gm;
    return;
  }
  const TexTaps<T> t = make_taps<T>(u, v, tp.Wt, tp.Ht);
  const int64_t o_nw = (int64_t)t.y0 * tp.Wt + t.x0, o_sw = (int64_t)t.y1 * tp.Wt + t.x0;
  const bool b_nw = t.vx0 && t.vy0, b_ne = t.vx1 && t.vy0, b_sw = t.vx0 && t.vy1, b_se = t.vx1 && t.vy1;
  const T fx = (T)t.x0, fy = (T)t.y0;
  const T ax = (fx + (T)1) - t.ix, bx = t.ix - fx, ay = (fy + (T)1) - t.iy, by = t.iy - fy;
  T gix = (T)0, giy = (T)0;
  for (int c = 0; c < tp.C; ++c) {
    const T* p = tex + c * plane;
    const T nw = b_nw ? p[o_nw] : (T)0, ne = b_ne ? p[o_nw + 1] : (T)0;
    const T sw = b_sw ? p[o_sw] : (T)0, se = b_se ? p[o_sw + 1] : (T)0;
    T gv = go[c];
    if (shade) {
      const T s = nw * t.wnw + ne * t.wne + sw * t.wsw + se * t.wse;
      const T pre = s * m;
      if (!(pre >= (T)0 && pre <= (T)1)) gv = (T)0;
      gm += gv * s;
      gv *= m;
    }
    if (gv == (T)0) continue;
    if (gtex) {
      T* q = gtex + c * plane;
      if (b_nw) atomicAdd(q + o_nw, gv * t.wnw);
      if (b_ne) atomicAdd(q + o_nw + 1, gv * t.wne);
      if (b_sw) atomicAdd(q + o_sw, gv * t.wsw);
      if (b_se) atomicAdd(q + o_sw + 1, gv * t.wse);
    }
    gix += gv * ((ne - nw) * ay + (se - sw) * by);
    giy += gv * ((sw - nw) * ax + (se - ne) * bx);
  }
  if (tp.g_uv) { tp.g_uv[g * 2] = gix * t.dix; tp.g_uv[g * 2 + 1] = giy * t.diy; }
  if (tp.g_mask) tp.g_mask[g] = gm;
}

}  // namespace dibr
