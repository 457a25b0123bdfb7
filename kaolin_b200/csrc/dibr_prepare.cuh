// dibr_prepare.cuh — the step immediately BEFORE the rasterization path:
// prepare_vertices of kaolin/render/mesh/utils.py:128-175, i.e.
//   rotate_translate_points / camera_transform        kaolin/render/camera/legacy.py:22-37, utils.py:163-167
//   perspective_camera                                 kaolin/render/camera/legacy.py:120-139
//   index_vertices_by_faces (x2)                       kaolin/ops/mesh/mesh.py:24-45
//   face_normals(unit=True)                            kaolin/ops/mesh/trianglemesh.py:313-336
// fused into ONE pass: thread = (view, face) gathers its three vertices, transforms and projects them and writes
// face_vertices_camera (B,F,3,3), face_vertices_image (B,F,3,2), face_normals (B,F,3).  The reference runs ~12
// torch kernels and materialises vertices_camera / vertices_image / two expanded gathers on the way.
// Backward: thread = (view, face) recomputes the camera-space triangle, folds the three upstream gradients
// through normalisation, cross product, perspective divide and rotation, and scatters into grad_vertices with
// red.global.add (vertices shared by all views when Bv == 1).
// Floating point: tolerance parity (1e-5 relative, fp32), not bit-exactness — the reference's own matmul has no
// specified summation order.
#pragma once
#include "dibr_common.cuh"

namespace dibr {

template <typename T>
struct PrepareParams {
  const T* vertices;        // (Bv,V,3), Bv = 1 (shared mesh) or B
  const int64_t* faces;     // (F,3)
  const T* proj;            // (3)
  const T* rot;             // (B,3,3) or nullptr
  const T* trans;           // (B,3)   or nullptr
  const T* transform;       // (B,4,3) or nullptr (replaces rot/trans)
  int B, Bv, V, F;
  // forward outputs
  T* face_cam;              // (B,F,3,3)
  T* face_img;              // (B,F,3,2)
  T* face_nrm;              // (B,F,3)
  // backward inputs (each may be nullptr) / output
  const T* g_cam;           // (B,F,3,3)
  const T* g_img;           // (B,F,3,2)
  const T* g_nrm;           // (B,F,3)
  T* g_vertices;            // (Bv,V,3), zero-filled by the host wrapper; nullptr = not wanted
  // optional camera gradients (zero-filled by the host wrapper; nullptr = not wanted): the reference is pure
  // torch and differentiates through every camera tensor (utils.py:160-172)
  T* g_proj;                // (3)
  T* g_rot;                 // (B,3,3)
  T* g_trans;               // (B,3)
  T* g_transform;           // (B,4,3)
};

template <typename T>
struct ViewCamera {         // cam_j = sum_k (p_k - t_k) * m[j][k]   (rot/trans)   or   sum_k p_k * m[j][k] + o[j]
  T m[3][3];
  T t[3];
  T o[3];
};

template <typename T>
__device__ __forceinline__ ViewCamera<T> load_camera(const PrepareParams<T>& pp, const int b) {
  ViewCamera<T> c;
  if (pp.transform) {                                  // padded_vertices @ camera_transform, utils.py:163-167
    const T* M = pp.transform + (int64_t)b * 12;       // (4,3): M[k][j]
#pragma unroll
    for (int j = 0; j < 3; ++j) {
#pragma unroll
      for (int k = 0; k < 3; ++k) c.m[j][k] = M[k * 3 + j];
      c.o[j] = M[9 + j];
      c.t[j] = (T)0;
    }
  } else {                                             // (points - trans) @ rot^T, legacy.py:35-37
    const T* R = pp.rot + (int64_t)b * 9;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
#pragma unroll
      for (int k = 0; k < 3; ++k) c.m[j][k] = R[j * 3 + k];
      c.o[j] = (T)0;
      c.t[j] = pp.trans[(int64_t)b * 3 + j];
    }
  }
  return c;
}

// Every operation of the forward is an explicit round-to-nearest intrinsic (never contracted or re-associated), so
// that the two kernels that project faces — prepare_vertices_fwd_kernel and the mesh form of setup_count_kernel
// (dibr_binning.cuh, dibr_rasterization_from_mesh) — produce the same bits: face_idx of the fused entry equals the
// unfused chain's.
template <typename T>
__device__ __forceinline__ void to_camera(const ViewCamera<T>& c, const T* p, T* q) {
  typedef RN<T> R;
  const T x = R::sub(p[0], c.t[0]), y = R::sub(p[1], c.t[1]), z = R::sub(p[2], c.t[2]);
#pragma unroll
  for (int j = 0; j < 3; ++j)
    q[j] = R::add(R::fma(z, c.m[j][2], R::fma(y, c.m[j][1], R::mul(x, c.m[j][0]))), c.o[j]);
}

template <typename T> __device__ __forceinline__ T sqrt_t(T v);
template <> __device__ __forceinline__ float sqrt_t<float>(float v) { return __fsqrt_rn(v); }
template <> __device__ __forceinline__ double sqrt_t<double>(double v) { return __dsqrt_rn(v); }

// One face of one view: camera-space vertices q, image coordinates img (legacy.py:136-137), unit normal
// (trianglemesh.py:328-334).  An out-of-range vertex index gives NaNs (never a read outside the buffer).
template <typename T>
__device__ __forceinline__ void project_face(const PrepareParams<T>& pp, const ViewCamera<T>& cam, const int b,
                                             const int f, T q[3][3], T img[6], T nrm[3]) {
  typedef RN<T> R;
  const T* vb = pp.vertices + (pp.Bv == 1 ? (int64_t)0 : (int64_t)b * pp.V * 3);
  const T p0 = pp.proj[0], p1 = pp.proj[1], p2 = pp.proj[2];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int64_t vi = pp.faces[(int64_t)f * 3 + k];
    if ((uint64_t)vi >= (uint64_t)pp.V) {
      q[k][0] = q[k][1] = q[k][2] = (T)NAN;
    } else {
      to_camera<T>(cam, vb + vi * 3, q[k]);
    }
    const T pz = R::mul(q[k][2], p2);
    img[k * 2 + 0] = R::div(R::mul(q[k][0], p0), pz);
    img[k * 2 + 1] = R::div(R::mul(q[k][1], p1), pz);
  }
  const T e0x = R::sub(q[1][0], q[0][0]), e0y = R::sub(q[1][1], q[0][1]), e0z = R::sub(q[1][2], q[0][2]);
  const T e1x = R::sub(q[2][0], q[0][0]), e1y = R::sub(q[2][1], q[0][1]), e1z = R::sub(q[2][2], q[0][2]);
  const T nx = R::fma(e0y, e1z, -R::mul(e0z, e1y)), ny = R::fma(e0z, e1x, -R::mul(e0x, e1z));
  const T nz = R::fma(e0x, e1y, -R::mul(e0y, e1x));
  const T len = R::add(sqrt_t<T>(R::fma(nz, nz, R::fma(ny, ny, R::mul(nx, nx)))), (T)1e-10);
  nrm[0] = R::div(nx, len); nrm[1] = R::div(ny, len); nrm[2] = R::div(nz, len);
}

template <typename T>
__global__ void __launch_bounds__(256) prepare_vertices_fwd_kernel(PrepareParams<T> pp) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (int64_t)pp.B * pp.F) return;
  const int b = (int)(g / pp.F), f = (int)(g % pp.F);
  const ViewCamera<T> cam = load_camera<T>(pp, b);
  T q[3][3], img[6], nrm[3];
  project_face<T>(pp, cam, b, f, q, img, nrm);
  T* oc = pp.face_cam + g * 9;
  T* oi = pp.face_img + g * 6;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    oc[k * 3 + 0] = q[k][0]; oc[k * 3 + 1] = q[k][1]; oc[k * 3 + 2] = q[k][2];
    oi[k * 2 + 0] = img[k * 2]; oi[k * 2 + 1] = img[k * 2 + 1];
  }
  T* on = pp.face_nrm + g * 3;
  on[0] = nrm[0]; on[1] = nrm[1]; on[2] = nrm[2];
}

template <typename T, bool CAMGRAD>
__device__ __forceinline__ void prepare_bwd_body(const PrepareParams<T>& pp, const int64_t g, const bool active,
                                                 const bool uniform, T (*acc)[3] = nullptr);

// Sum of `v` over the lanes of a warp whose active lanes all belong to view `b`; one atomic by lane 0.  Warps that
// straddle two views (or hold inactive lanes) fall back to one atomic per lane.
template <typename T>
__device__ __forceinline__ void warp_accumulate(T* dst, T v, const bool uniform) {
  if (uniform) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v != (T)0) atomicAdd(dst, v);
  } else if (v != (T)0) {
    atomicAdd(dst, v);
  }
}

template <typename T, bool CAMGRAD>
__global__ void __launch_bounds__(256) prepare_vertices_bwd_kernel(PrepareParams<T> pp) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool in_range = g < (int64_t)pp.B * pp.F;
  if (!CAMGRAD && !in_range) return;
  if (CAMGRAD) {                     // the camera sums are warp-collective: out-of-range lanes contribute zeros
    const int64_t gl = in_range ? g : (int64_t)pp.B * pp.F - 1;
    const int bl = (int)(gl / pp.F);
    const int b0 = __shfl_sync(0xffffffffu, bl, 0);
    const bool uniform = __all_sync(0xffffffffu, bl == b0);
    prepare_bwd_body<T, true>(pp, gl, in_range, uniform);
  } else {
    prepare_bwd_body<T, false>(pp, g, true, false);
  }
}

// One mesh seen by B cameras (Bv == 1), no camera gradients: thread = (face, chunk of views).  The face's gradient is
// summed over the chunk's views in registers and leaves as 9 atomics per chunk instead of 9 per view — every view
// scatters into the SAME V x 3 words, and the L2 atomic unit serialises per address (c4: 96 -> 24 / 12 atomics per
// coordinate at 16 / 128 cameras).  grid (ceil(F / 256), ceil(B / views_per_thread)).
template <typename T>
__global__ void __launch_bounds__(256) prepare_vertices_bwd_shared_kernel(PrepareParams<T> pp,
                                                                          const int views_per_thread) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= pp.F) return;
  T acc[3][3];
#pragma unroll
  for (int k = 0; k < 3; ++k) acc[k][0] = acc[k][1] = acc[k][2] = (T)0;
  const int b0 = blockIdx.y * views_per_thread;
  const int b1 = b0 + views_per_thread < pp.B ? b0 + views_per_thread : pp.B;
  for (int b = b0; b < b1; ++b) prepare_bwd_body<T, false>(pp, (int64_t)b * pp.F + f, true, false, acc);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int64_t vi = pp.faces[(int64_t)f * 3 + k];
    if ((uint64_t)vi >= (uint64_t)pp.V) return;        // (the body added nothing for this face)
#pragma unroll
    for (int c = 0; c < 3; ++c)
      if (acc[k][c] != (T)0) atomicAdd(pp.g_vertices + vi * 3 + c, acc[k][c]);
  }
}

template <typename T, bool CAMGRAD>
__device__ __forceinline__ void prepare_bwd_body(const PrepareParams<T>& pp, const int64_t g, const bool active,
                                                 const bool uniform, T (*acc)[3]) {
  const int b = (int)(g / pp.F), f = (int)(g % pp.F);
  const ViewCamera<T> cam = load_camera<T>(pp, b);
  const int64_t voff = pp.Bv == 1 ? (int64_t)0 : (int64_t)b * pp.V * 3;
  const T* vb = pp.vertices + voff;
  const T p0 = pp.proj[0], p1 = pp.proj[1], p2 = pp.proj[2];
  int64_t vi[3];
  T q[3][3], gq[3][3];
  T gp0 = 0, gp1 = 0, gp2 = 0;
  bool ok = active;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    vi[k] = pp.faces[(int64_t)f * 3 + k];
    if ((uint64_t)vi[k] >= (uint64_t)pp.V) { ok = false; vi[k] = 0; }  // forward wrote NaN for this face: no gradient
  }
  if (!CAMGRAD && !ok) return;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    to_camera<T>(cam, vb + vi[k] * 3, q[k]);
#pragma unroll
    for (int j = 0; j < 3; ++j) gq[k][j] = (ok && pp.g_cam) ? pp.g_cam[g * 9 + k * 3 + j] : (T)0;
  }
  if (ok && pp.g_img) {               // img_x = (x p0) / (z p2),  img_y = (y p1) / (z p2)
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const T gx = pp.g_img[g * 6 + k * 2], gy = pp.g_img[g * 6 + k * 2 + 1];
      const T pz = q[k][2] * p2;
      const T ix = (q[k][0] * p0) / pz, iy = (q[k][1] * p1) / pz;
      gq[k][0] += gx * p0 / pz;
      gq[k][1] += gy * p1 / pz;
      gq[k][2] -= (gx * ix + gy * iy) / q[k][2];
      if (CAMGRAD) {                  // d ix / d p0 = x / pz,  d ix / d p2 = -ix / p2
        gp0 += gx * q[k][0] / pz;
        gp1 += gy * q[k][1] / pz;
        gp2 -= (gx * ix + gy * iy) / p2;
      }
    }
  }
  if (ok && pp.g_nrm) {               // n = c / (|c| + 1e-10), c = e0 x e1
    const T e0x = q[1][0] - q[0][0], e0y = q[1][1] - q[0][1], e0z = q[1][2] - q[0][2];
    const T e1x = q[2][0] - q[0][0], e1y = q[2][1] - q[0][1], e1z = q[2][2] - q[0][2];
    const T cx = e0y * e1z - e0z * e1y, cy = e0z * e1x - e0x * e1z, cz = e0x * e1y - e0y * e1x;
    const T L = sqrt_t<T>(cx * cx + cy * cy + cz * cz);
    const T Le = L + (T)1e-10;
    const T gx = pp.g_nrm[g * 3], gy = pp.g_nrm[g * 3 + 1], gz = pp.g_nrm[g * 3 + 2];
    const T dot = cx * gx + cy * gy + cz * gz;
    const T s = L > (T)0 ? dot / (L * Le * Le) : (T)0;           // torch's norm backward is 0 at 0
    const T hx = gx / Le - cx * s, hy = gy / Le - cy * s, hz = gz / Le - cz * s;     // dL/dc
    // c = e0 x e1:  dL/de0 = e1 x h,  dL/de1 = h x e0
    const T a0 = e1y * hz - e1z * hy, a1 = e1z * hx - e1x * hz, a2 = e1x * hy - e1y * hx;
    const T b0 = hy * e0z - hz * e0y, b1 = hz * e0x - hx * e0z, b2 = hx * e0y - hy * e0x;
    gq[1][0] += a0; gq[1][1] += a1; gq[1][2] += a2;
    gq[2][0] += b0; gq[2][1] += b1; gq[2][2] += b2;
    gq[0][0] -= a0 + b0; gq[0][1] -= a1 + b1; gq[0][2] -= a2 + b2;
  }
  T gt[3] = {0, 0, 0};                // sum over the face's vertices of dL/dp (rot/trans form: dL/dtrans = -that)
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {     // cam_j = sum_c (p_c - t_c) m[j][c]  =>  dL/dp_c = sum_j gq_j m[j][c]
      const T v = gq[k][0] * cam.m[0][c] + gq[k][1] * cam.m[1][c] + gq[k][2] * cam.m[2][c];
      if (ok && pp.g_vertices) {
        if (acc) acc[k][c] += v;                       // summed over a chunk of views by the caller
        else atomicAdd(pp.g_vertices + voff + vi[k] * 3 + c, v);
      }
      gt[c] += v;
    }
  }
  if (!CAMGRAD) return;
  if (pp.g_proj) {
    warp_accumulate<T>(pp.g_proj + 0, gp0, true);
    warp_accumulate<T>(pp.g_proj + 1, gp1, true);
    warp_accumulate<T>(pp.g_proj + 2, gp2, true);
  }
  if (pp.g_trans) {
#pragma unroll
    for (int c = 0; c < 3; ++c) warp_accumulate<T>(pp.g_trans + (int64_t)b * 3 + c, ok ? -gt[c] : (T)0, uniform);
  }
  if (pp.g_rot || pp.g_transform) {
    // dL/dm[j][c] = sum over the face's vertices of gq[.][j] * (p_c - t_c)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        T v = 0;
#pragma unroll
        for (int k = 0; k < 3; ++k) v += gq[k][j] * (vb[vi[k] * 3 + c] - cam.t[c]);
        v = ok ? v : (T)0;
        if (pp.g_rot) warp_accumulate<T>(pp.g_rot + (int64_t)b * 9 + j * 3 + c, v, uniform);
        else warp_accumulate<T>(pp.g_transform + (int64_t)b * 12 + c * 3 + j, v, uniform);      // M[c][j]
      }
      if (pp.g_transform) {           // cam_j += M[3][j]
        const T v = ok ? gq[0][j] + gq[1][j] + gq[2][j] : (T)0;
        warp_accumulate<T>(pp.g_transform + (int64_t)b * 12 + 9 + j, v, uniform);
      }
    }
  }
}

}  // namespace dibr
