// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement (C++17, OpenMP) of the four reference kernels on the DIB-R hot path (and of the two DefTet
// sparse-renderer kernels).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library; kaolin_b200/ never does (the product path fails loudly without its CUDA .so).
//
// Parity status: PINNED.  Checked against (a) the reference's golden .pt vectors
// (tests/samples/dibr/{simple,sphere}, converted by tests/golden/make_golden.py), (b) outputs of the
// reference's own pure-torch `_naive_deftet_sparse_render` generated in the build container, and
// (c) on the GPU box, the reference's unmodified CUDA kernels built as oracle/_ref/ref_C*.so.
//
// Every function cites the reference lines it follows (paths relative to /root/reference/).
// The per-(pixel, face) floating-point sequences are written with explicit fma/mul/add so that
// they reproduce, operation for operation, what nvcc (default -fmad=true, -prec-div=true) emits for
// the reference source on sm_100a; the sequences were read from the reference's SASS
// (cuobjdump of kaolin/csrc/render/mesh/{rasterization,dibr_soft_mask}_cuda.cu).
// This file must be compiled with -ffp-contract=off so the compiler adds no contraction of its own.

#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>
#include <algorithm>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

template <typename T> inline T rfma(T a, T b, T c) { return std::fma(a, b, c); }

// pixel centre — kaolin/csrc/render/mesh/rasterization_cuda.cu:85-86 (also dibr_soft_mask_cuda.cu:75-76,
// :259-260).  `multiplier` is a C float and width/height ints, so the value is computed in fp32
// (one IEEE division, one multiplication) even for the double instantiation, then widened.
inline float pix_x(float multiplier, int width, int wididx) {
  return multiplier / (float)width * (float)(2 * wididx + 1 - width);
}
inline float pix_y(float multiplier, int height, int heiidx) {
  return multiplier / (float)height * (float)(height - 2 * heiidx - 1);
}

// ---------------------------------------------------------------------------------------------
// rasterization_cuda.cu:43-192  packed_rasterize_forward_cuda_kernel
// ---------------------------------------------------------------------------------------------
template <typename T>
void packed_rasterize_forward(const T* z, const T* img, const T* bbox, const T* feat,
                              const int64_t* first_idx, int64_t* out_idx, T* out_w, T* out_feat,
                              int B, int H, int W, int C, float multiplier, float eps,
                              int row_begin, int row_end) {
  // outputs are pre-initialised by the caller as in rasterization.cpp:87-93 (idx=-1, weights=0, feat=0)
  if (row_end < 0) row_end = H;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
  for (int b = 0; b < B; ++b) {
    for (int j = row_begin; j < row_end; ++j) {
      const int64_t f0 = first_idx[b], f1 = first_idx[b + 1];   // :71-72
      for (int i = 0; i < W; ++i) {
        const T x0 = (T)pix_x(multiplier, W, i);
        const T y0 = (T)pix_y(multiplier, H, j);
        T max_z0 = -std::numeric_limits<T>::infinity();        // :73
        int64_t best = -1;
        T bw0 = 0, bw1 = 0, bw2 = 0;
        for (int64_t f = f0; f < f1; ++f) {
          const T* bb = bbox + f * 4;
          if (x0 < bb[0] || x0 >= bb[2] || y0 < bb[1] || y0 >= bb[3]) continue;   // :115
          const T* p = img + f * 6;
          const T aex = p[0] - x0, aey = p[1] - y0;               // :131-136
          const T bex = p[2] - x0, bey = p[3] - y0;
          const T cex = p[4] - x0, cey = p[5] - y0;
          // :137-139 — nvcc fuses the first product of each difference, rounds the second
          T w0 = rfma(bex, cey, -(bey * cex));
          T w1 = rfma(cex, aey, -(cey * aex));
          T w2 = rfma(aex, bey, -(aey * bex));
          T norm = (w0 + w1) + w2;                                 // :140
          norm = (T)((double)norm + std::copysign((double)eps, (double)norm));   // :141-142
          w0 /= norm; w1 /= norm; w2 /= norm;                      // :143-145
          if (w0 < 0. || w1 < 0. || w2 < 0.) continue;             // :148
          const T* zz = z + f * 3;
          const T z0 = rfma(zz[2], w2, rfma(zz[0], w0, zz[1] * w1));   // :159 (SASS order)
          if (z0 <= max_z0) continue;                              // :162 (NaN is accepted)
          max_z0 = z0; best = f; bw0 = w0; bw1 = w1; bw2 = w2;     // :165-169
        }
        if (best > -1) {                                           // :172-189
          const int64_t pix = ((int64_t)b * H + j) * W + i;
          out_idx[pix] = best - f0;
          out_w[pix * 3 + 0] = bw0; out_w[pix * 3 + 1] = bw1; out_w[pix * 3 + 2] = bw2;
          const T* ff = feat + best * 3 * C;
          for (int d = 0; d < C; ++d)
            out_feat[pix * C + d] = rfma(ff[2 * C + d], bw2, rfma(ff[d], bw0, ff[C + d] * bw1));
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// rasterization_cuda.cu:238-402  rasterize_backward_cuda_kernel
// (gradient accumulation order is unspecified in the reference — atomicAdd — so no rounding
//  sequence is pinned here; tolerance oracle, 1e-4 relative)
// ---------------------------------------------------------------------------------------------
template <typename T>
void rasterize_backward(const T* grad, const int64_t* idx, const T* wts, const T* img, const T* feat,
                        T* grad_img, T* grad_feat, int B, int H, int W, int F, int C, float eps,
                        int row_begin, int row_end) {
  if (row_end < 0) row_end = H;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
  for (int b = 0; b < B; ++b) {
    for (int j = row_begin; j < row_end; ++j) {
      for (int i = 0; i < W; ++i) {
        const int64_t pix = ((int64_t)b * H + j) * W + i;
        const int64_t fi = idx[pix];
        if (fi < 0) continue;                                      // :266
        const int64_t tf = (int64_t)b * F + fi;
        const T* g = grad + pix * C;
        const T* w = wts + pix * 3;
        T* gf = grad_feat + tf * 3 * C;
        for (int k = 0; k < 3; ++k)                                // :273-285
          for (int d = 0; d < C; ++d) {
            const T v = g[d] * w[k];
#pragma omp atomic
            gf[k * C + d] += v;
          }
        const T* p = img + tf * 6;                                 // :291-296
        const T ax = p[0], ay = p[1], bx = p[2], by = p[3], cx = p[4], cy = p[5];
        const T aw = w[0], bw = w[1], cw = w[2];
        const T x0 = aw * ax + bw * bx + cw * cx;                  // :302-303
        const T y0 = aw * ay + bw * by + cw * cy;
        const T m = bx - ax, pp = by - ay, n = cx - ax, q = cy - ay, s = x0 - ax, t = y0 - ay;
        const T k1 = s * q - n * t;                                // :319-322
        const T k2 = m * t - s * pp;
        T k3 = m * q - n * pp;
        k3 = (T)((double)k3 + std::copysign((double)eps, (double)k3));
        // :324-374 closed-form numerators of d w1, d w2 (the common 1/k3^2 is applied in dldI)
        const T dw1dm = -q * k1, dw1dn = -t * k3 + pp * k1, dw1dp = n * k1, dw1dq = s * k3 - m * k1;
        const T dw1ds = q * k3, dw1dt = -n * k3;
        const T dw2dm = t * k3 - q * k2, dw2dn = pp * k2, dw2dp = -s * k3 + n * k2, dw2dq = -m * k2;
        const T dw2ds = -pp * k3, dw2dt = m * k3;
        const T dw1dax = -(dw1dm + dw1dn + dw1ds), dw1day = -(dw1dp + dw1dq + dw1dt);
        const T dw2dax = -(dw2dm + dw2dn + dw2ds), dw2day = -(dw2dp + dw2dq + dw2dt);
        T acc[6] = {0, 0, 0, 0, 0, 0};
        const T* ff = feat + tf * 3 * C;
        for (int d = 0; d < C; ++d) {                              // :376-399
          const T c0 = ff[d], c1 = ff[C + d], c2 = ff[2 * C + d];
          const T dldI = g[d] / (k3 * k3);
          acc[0] += dldI * ((c1 - c0) * dw1dax + (c2 - c0) * dw2dax);
          acc[1] += dldI * ((c1 - c0) * dw1day + (c2 - c0) * dw2day);
          acc[2] += dldI * ((c1 - c0) * dw1dm + (c2 - c0) * dw2dm);
          acc[3] += dldI * ((c1 - c0) * dw1dp + (c2 - c0) * dw2dp);
          acc[4] += dldI * ((c1 - c0) * dw1dn + (c2 - c0) * dw2dn);
          acc[5] += dldI * ((c1 - c0) * dw1dq + (c2 - c0) * dw2dq);
        }
        T* gi = grad_img + tf * 6;
        for (int k = 0; k < 6; ++k) {
#pragma omp atomic
          gi[k] += acc[k];
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// dibr_soft_mask_cuda.cu:27-184  dibr_soft_mask_forward_cuda_kernel
// `#define EPS 1e-7` (:23) is a double literal: (down + EPS) and the three quotients per edge are
// evaluated in double and rounded back to scalar_t, as are (1.0 - prob) and the running product.
// ---------------------------------------------------------------------------------------------
template <typename T> inline T exp_neg(T z);
template <> inline float exp_neg<float>(float z) { return std::exp(-z); }   // expf; CUDA's expf may differ by ~1 ulp
template <> inline double exp_neg<double>(double z) { return std::exp(-z); }

template <typename T>
inline void soft_face_distance(const T* p, T x0, T y0, float multiplier, T& dissquare, int& edgeid) {
  const double EPS = 1e-7;
  T pdis[6];
  for (int i = 0; i < 3; ++i) {                                    // :103-140
    const T x1 = p[i * 2], y1 = p[i * 2 + 1];
    const int i2 = (i + 1) % 3;
    const T x2 = p[i2 * 2], y2 = p[i2 * 2 + 1];
    const T A = y2 - y1, Bc = x1 - x2;                             // :114-115
    const T Cc = rfma(x2, y1, -(x1 * y2));                         // :116
    const T AA = A * A, BB = Bc * Bc, AB = A * Bc;
    const T up = Cc + rfma(y0, Bc, x0 * A);                        // :122
    const T down = AA + BB;                                        // :123
    T x3 = rfma(-A, Cc, rfma(x0, BB, -(y0 * AB)));                 // :126
    T y3 = rfma(-Bc, Cc, rfma(y0, AA, -(x0 * AB)));                // :127
    const double den = (double)down + EPS;
    x3 = (T)((double)x3 / den);                                    // :128-129
    y3 = (T)((double)y3 / den);
    const T direct = rfma(x3 - x1, x3 - x2, (y3 - y1) * (y3 - y2));   // :131
    if (direct > 0) {
      pdis[i] = (T)(4 * multiplier * multiplier);                  // :135 (fp32 arithmetic)
    } else {
      pdis[i] = (T)((double)(up * up) / den);                      // :138
    }
  }
  for (int i = 0; i < 3; ++i) {                                    // :144-149
    const T dx = x0 - p[i * 2], dy = y0 - p[i * 2 + 1];
    pdis[i + 3] = rfma(dx, dx, dy * dy);
  }
  edgeid = 0; dissquare = pdis[0];                                 // :151-159 (first minimum wins)
  for (int i = 1; i < 6; ++i)
    if (dissquare > pdis[i]) { dissquare = pdis[i]; edgeid = i; }
}

template <typename T>
void dibr_soft_mask_forward(const T* img, const T* bbox, const int64_t* sel, T* prob_out,
                            int64_t* cidx_out, uint8_t* ctype_out, T* soft, int B, int H, int W,
                            int F, int knum, float sigmainv, float multiplier,
                            int row_begin, int row_end) {
  // prob/cidx/ctype may be null (then only `soft` is produced). Caller pre-initialises them as
  // dibr_soft_mask.cpp:84-97 (idx=-1, prob=0, type=0, soft=0).
  if (row_end < 0) row_end = H;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
  for (int b = 0; b < B; ++b) {
    for (int j = row_begin; j < row_end; ++j) {
      std::vector<T> probs((size_t)knum);
      for (int i = 0; i < W; ++i) {
        const int64_t pix = ((int64_t)b * H + j) * W + i;
        if (sel[pix] >= 0) { soft[pix] = (T)1.0; continue; }      // :68-70
        const T x0 = (T)pix_x(multiplier, W, i);
        const T y0 = (T)pix_y(multiplier, H, j);
        int kid = 0;
        for (int f = 0; f < F; ++f) {                              // :80
          const int64_t s1 = (int64_t)b * F + f;
          const T* bb = bbox + s1 * 4;
          if (x0 < bb[0] || x0 >= bb[2] || y0 < bb[1] || y0 >= bb[3]) continue;   // :95
          T dissquare; int edgeid;
          soft_face_distance<T>(img + s1 * 6, x0, y0, multiplier, dissquare, edgeid);
          const T zz = (T)sigmainv * dissquare / (T)multiplier / (T)multiplier;   // :161
          const T prob = exp_neg<T>(zz);                           // :163
          probs[kid] = prob;
          if (prob_out) {
            prob_out[pix * knum + kid] = prob;                     // :165-167
            cidx_out[pix * knum + kid] = f;
            ctype_out[pix * knum + kid] = (uint8_t)(edgeid + 1);
          }
          if (++kid >= knum) break;                                // :168-171
        }
        T allprob = (T)1.0;                                        // :174-182
        for (int k = 0; k < kid; ++k) allprob = (T)((double)allprob * (1.0 - (double)probs[k]));
        soft[pix] = (T)(1.0 - (double)allprob);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// dibr_soft_mask_cuda.cu:230-353  dibr_soft_mask_backward_cuda_kernel
// ---------------------------------------------------------------------------------------------
template <typename T>
void dibr_soft_mask_backward(const T* gsoft, const T* soft, const int64_t* sel, const T* prob_in,
                             const int64_t* cidx, const uint8_t* ctype, const T* img, T* grad_img,
                             int B, int H, int W, int F, int knum, float sigmainv, float multiplier,
                             int row_begin, int row_end) {
  const double EPS = 1e-7;
  if (row_end < 0) row_end = H;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
  for (int b = 0; b < B; ++b) {
    for (int j = row_begin; j < row_end; ++j) {
      for (int i = 0; i < W; ++i) {
        const int64_t pix = ((int64_t)b * H + j) * W + i;
        if (sel[pix] >= 0) continue;                               // :266
        const T x0 = (T)pix_x(multiplier, W, i);
        const T y0 = (T)pix_y(multiplier, H, j);
        const T dLdp = gsoft[pix], allprob = soft[pix];
        for (int k = 0; k < knum; ++k) {                           // :271-276
          const int64_t f = cidx[pix * knum + k];
          if (f < 0) break;
          const int64_t s6 = ((int64_t)b * F + f) * 6;
          const T prob = prob_in[pix * knum + k];
          const T dLdz = (T)(-1.0 * sigmainv * dLdp * (1.0 - allprob) / (1.0 - prob + EPS) * prob);   // :283
          const int edgeid = (int)ctype[pix * knum + k] - 1;
          auto add = [&](int64_t off, T v) {
#pragma omp atomic
            grad_img[off] += v;
          };
          if (edgeid >= 3) {                                       // :289-302
            const int64_t ps = s6 + (edgeid - 3) * 2;
            const T x1 = img[ps], y1 = img[ps + 1];
            add(ps, dLdz * 2 * (x1 - x0) / multiplier);
            add(ps + 1, dLdz * 2 * (y1 - y0) / multiplier);
          } else {                                                 // :304-347
            const int64_t ps = s6 + edgeid * 2, ps2 = s6 + ((edgeid + 1) % 3) * 2;
            const T x1 = img[ps], y1 = img[ps + 1], x2 = img[ps2], y2 = img[ps2 + 1];
            const T A = y2 - y1, Bc = x1 - x2, Cc = x2 * y1 - x1 * y2;
            const T up = A * x0 + Bc * y0 + Cc;
            const T down = A * A + Bc * Bc;
            const T dissquare = (T)(up * up / (down + EPS));
            const T dzdA = (T)(2 * (x0 * up - dissquare * A) / (down + EPS));
            const T dzdB = (T)(2 * (y0 * up - dissquare * Bc) / (down + EPS));
            const T dzdC = (T)(2 * up / (down + EPS));
            add(ps, dLdz * (dzdB - y2 * dzdC) / multiplier);
            add(ps + 1, dLdz * (x2 * dzdC - dzdA) / multiplier);
            add(ps2, dLdz * (y1 * dzdC - dzdB) / multiplier);
            add(ps2 + 1, dLdz * (dzdA - x1 * dzdC) / multiplier);
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// deftet_cuda.cu:31-190  deftet_sparse_render_forward_cuda_kernel   
// Per pixel: the first `knum` faces IN FACE ORDER whose half-open bbox contains the pixel, whose
// barycentric weights are all >= 0 and whose interpolated depth lies in [min_depth, max_depth), written
// in insertion order (the depth sort is torch glue, deftet.py:291-296).  Outputs pre-initialised by the
// caller as in deftet.cpp:87-92 (idx = -1, depth = -inf, w0 = w1 = 0).
// Contractions read from the reference's sm_100a SASS: w = fma(p, q, -(r*s)); norm = (w0+w1)+w2;
// depth = fma(w2, cz, fma(w0, az, w1*bz)); one correctly rounded division per weight.
// ---------------------------------------------------------------------------------------------
template <typename T>
void deftet_forward(const T* z, const T* img, const T* bbox, const T* pix, const T* ranges, int64_t* out_idx,
                    T* out_depth, T* out_w0, T* out_w1, int B, int F, int P, int knum, float eps) {
#pragma omp parallel for collapse(2) schedule(dynamic, 16)
  for (int b = 0; b < B; ++b) {
    for (int p = 0; p < P; ++p) {
      const int64_t m = (int64_t)b * P + p;
      const T x0 = pix[m * 2], y0 = pix[m * 2 + 1];
      const T min_depth = ranges[m * 2], max_depth = ranges[m * 2 + 1];
      int n = 0;
      for (int f = 0; f < F && n < knum; ++f) {
        const int64_t g = (int64_t)b * F + f;
        const T* bb = bbox + g * 4;
        if (!(x0 >= bb[0] && x0 < bb[2] && y0 >= bb[1] && y0 < bb[3])) continue;      // :124
        const T* v = img + g * 6;
        const T aex = v[0] - x0, aey = v[1] - y0, bex = v[2] - x0, bey = v[3] - y0, cex = v[4] - x0,
                cey = v[5] - y0;                                                        // :133-138
        const T w0n = rfma(bex, cey, -(bey * cex));                                     // :139
        const T w1n = rfma(cex, aey, -(cey * aex));                                     // :140
        const T w2n = rfma(aex, bey, -(aey * bex));                                     // :141
        const T norm = (w0n + w1n) + w2n;                                               // :142
        const T den = norm + (T)std::copysign(eps, (float)norm);                        // :143-146 (copysignf)
        const T w0 = w0n / den, w1 = w1n / den, w2 = w2n / den;
        if (!(w0 >= 0. && w1 >= 0. && w2 >= 0.)) continue;                              // :149
        const T* zz = z + g * 3;
        const T depth = rfma(w2, zz[2], rfma(w0, zz[0], w1 * zz[1]));                   // :158
        if (!(depth < max_depth && depth >= min_depth)) continue;                       // :160
        const int64_t o = m * knum + n;
        out_idx[o] = f; out_w0[o] = w0; out_w1[o] = w1; out_depth[o] = depth;           // :181-185
        ++n;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// deftet_cuda.cu:240-410  deftet_sparse_render_backward_cuda_kernel.  One (pixel, k) entry at a time; the
// gradients are accumulated serially here (atomicAdd order is unspecified in the reference), tolerance parity.
// ---------------------------------------------------------------------------------------------
template <typename T>
void deftet_backward(const T* grad, const int64_t* idx, const T* wts, const T* img, const T* feat, T* grad_img,
                     T* grad_feat, int B, int F, int P, int knum, int C, float eps) {
  for (int64_t e = 0; e < (int64_t)B * P * knum; ++e) {
    const int64_t f = idx[e];
    if (f < 0) continue;
    const int b = (int)(e / ((int64_t)P * knum));
    const int64_t g = (int64_t)b * F + f;
    const T* w = wts + e * 3;
    const T* go = grad + e * C;
    const T* ff = feat + g * 3 * C;
    for (int ii = 0; ii < 3; ++ii)
      for (int c = 0; c < C; ++c) grad_feat[g * 3 * C + ii * C + c] += go[c] * w[ii];   // :277-288
    const T* v = img + g * 6;
    const T ax = v[0], ay = v[1], bx = v[2], by = v[3], cx = v[4], cy = v[5];
    const T x0 = w[0] * ax + w[1] * bx + w[2] * cx, y0 = w[0] * ay + w[1] * by + w[2] * cy;   // :305-306
    const T m = bx - ax, p = by - ay, n = cx - ax, q = cy - ay, s = x0 - ax, t = y0 - ay;
    const T k1 = s * q - n * t, k2 = m * t - s * p;
    T k3 = m * q - n * p;
    k3 += (T)std::copysign((double)eps, (double)k3);                                     // :326-328
    const T dw1dm = -q * k1, dw1dn = -t * k3 + p * k1, dw1dp = n * k1, dw1dq = s * k3 - m * k1,
            dw1ds = q * k3, dw1dt = -n * k3;                                             // :352-357
    const T dw2dm = t * k3 - q * k2, dw2dn = p * k2, dw2dp = -s * k3 + n * k2, dw2dq = -m * k2,
            dw2ds = -p * k3, dw2dt = m * k3;                                             // :359-364
    const T dw1dax = -(dw1dm + dw1dn + dw1ds), dw1day = -(dw1dp + dw1dq + dw1dt);
    const T dw2dax = -(dw2dm + dw2dn + dw2ds), dw2day = -(dw2dp + dw2dq + dw2dt);
    for (int c = 0; c < C; ++c) {
      const T c0 = ff[c], c1 = ff[C + c], c2 = ff[2 * C + c];
      const T dldI = go[c] / (k3 * k3);                                                  // :394
      T* gi = grad_img + g * 6;
      gi[0] += dldI * ((c1 - c0) * dw1dax + (c2 - c0) * dw2dax);
      gi[1] += dldI * ((c1 - c0) * dw1day + (c2 - c0) * dw2day);
      gi[2] += dldI * ((c1 - c0) * dw1dm + (c2 - c0) * dw2dm);
      gi[3] += dldI * ((c1 - c0) * dw1dp + (c2 - c0) * dw2dp);
      gi[4] += dldI * ((c1 - c0) * dw1dn + (c2 - c0) * dw2dn);
      gi[5] += dldI * ((c1 - c0) * dw1dq + (c2 - c0) * dw2dq);
    }
  }
}

}  // namespace

#define ORACLE_API extern "C" __attribute__((visibility("default")))

#define DEFINE_ALL(SUF, T)                                                                          \
  ORACLE_API void oracle_packed_rasterize_forward_##SUF(                                           \
      const T* z, const T* img, const T* bbox, const T* feat, const int64_t* first_idx,             \
      int64_t* out_idx, T* out_w, T* out_feat, int B, int H, int W, int C, float multiplier,        \
      float eps, int row_begin, int row_end) {                                                      \
    packed_rasterize_forward<T>(z, img, bbox, feat, first_idx, out_idx, out_w, out_feat, B, H, W,   \
                                C, multiplier, eps, row_begin, row_end);                            \
  }                                                                                                 \
  ORACLE_API void oracle_rasterize_backward_##SUF(                                                  \
      const T* grad, const int64_t* idx, const T* wts, const T* img, const T* feat, T* grad_img,    \
      T* grad_feat, int B, int H, int W, int F, int C, float eps, int row_begin, int row_end) {     \
    rasterize_backward<T>(grad, idx, wts, img, feat, grad_img, grad_feat, B, H, W, F, C, eps,       \
                          row_begin, row_end);                                                      \
  }                                                                                                 \
  ORACLE_API void oracle_dibr_soft_mask_forward_##SUF(                                              \
      const T* img, const T* bbox, const int64_t* sel, T* prob, int64_t* cidx, uint8_t* ctype,      \
      T* soft, int B, int H, int W, int F, int knum, float sigmainv, float multiplier,              \
      int row_begin, int row_end) {                                                                 \
    dibr_soft_mask_forward<T>(img, bbox, sel, prob, cidx, ctype, soft, B, H, W, F, knum, sigmainv,  \
                              multiplier, row_begin, row_end);                                      \
  }                                                                                                 \
  ORACLE_API void oracle_dibr_soft_mask_backward_##SUF(                                             \
      const T* gsoft, const T* soft, const int64_t* sel, const T* prob, const int64_t* cidx,        \
      const uint8_t* ctype, const T* img, T* grad_img, int B, int H, int W, int F, int knum,        \
      float sigmainv, float multiplier, int row_begin, int row_end) {                               \
    dibr_soft_mask_backward<T>(gsoft, soft, sel, prob, cidx, ctype, img, grad_img, B, H, W, F,      \
                               knum, sigmainv, multiplier, row_begin, row_end);                     \
  }                                                                                                 \
  ORACLE_API void oracle_deftet_forward_##SUF(                                                      \
      const T* z, const T* img, const T* bbox, const T* pix, const T* ranges, int64_t* out_idx,     \
      T* out_depth, T* out_w0, T* out_w1, int B, int F, int P, int knum, float eps) {               \
    deftet_forward<T>(z, img, bbox, pix, ranges, out_idx, out_depth, out_w0, out_w1, B, F, P, knum, \
                      eps);                                                                         \
  }                                                                                                 \
  ORACLE_API void oracle_deftet_backward_##SUF(                                                     \
      const T* grad, const int64_t* idx, const T* wts, const T* img, const T* feat, T* grad_img,    \
      T* grad_feat, int B, int F, int P, int knum, int C, float eps) {                              \
    deftet_backward<T>(grad, idx, wts, img, feat, grad_img, grad_feat, B, F, P, knum, C, eps);      \
  }

DEFINE_ALL(f32, float)
DEFINE_ALL(f64, double)

ORACLE_API int oracle_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
ORACLE_API void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}

// float(double(n) + copysign(eps, n)) == n +_f32 copysignf(eps, n): exhaustive-ish self check used by
// tests/test_oracle.py to justify the fp32 shortcut taken in the CUDA kernel.
ORACLE_API int64_t oracle_check_eps_add_equivalence(const float* vals, int64_t n, float eps) {
  int64_t bad = 0;
  for (int64_t i = 0; i < n; ++i) {
    const float v = vals[i];
    const float a = (float)((double)v + std::copysign((double)eps, (double)v));
    const float b = v + std::copysign(eps, v);
    if (std::memcmp(&a, &b, 4) != 0 && !(std::isnan(a) && std::isnan(b))) ++bad;
  }
  return bad;
}
