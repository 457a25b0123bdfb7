/*
 * dibr_b200.h — C ABI of libdibr_b200.so: the B200 (sm_100a) implementation of Kaolin's DIB-R
 * differentiable rasterization hot path.
 *
 * This is the drop-in boundary.  It replaces the four pybind functions the reference exposes as
 * kaolin._C.render.mesh.* (kaolin/csrc/bindings.cpp:75-80, C++ signatures in
 * kaolin/csrc/render/mesh/rasterization.h:23-41 and dibr_soft_mask.h:23-40) and, in the `_fused`
 * entry points, also the torch glue the reference runs around them (kaolin/render/mesh/
 * rasterization.py:299-363, dibr.py:31-39).
 *
 * Conventions
 *   - extern "C", plain device pointers + sizes, no torch types.  All pointers are DEVICE pointers
 *     on the CURRENT device unless stated otherwise; tensors are dense row-major ("contiguous").
 *   - `stream` is a cudaStream_t (passed as void*); every call is asynchronous w.r.t. the host and
 *     enqueues work only on that stream.  The library never synchronises, never allocates device
 *     memory and keeps no mutable global state: it is re-entrant (forward runs on the Python
 *     thread, backward on the autograd thread — SURVEY.md §8b).
 *   - The caller allocates every output and the scratch workspace (`ws`, at least
 *     dibr_workspace_bytes(...) bytes, 256-byte aligned).  Outputs need NOT be pre-initialised:
 *     the kernels write every element (the reference's at::zeros/at::full fills,
 *     rasterization.cpp:87-93, dibr_soft_mask.cpp:84-97, are fused away).  Gradient outputs are
 *     zero-filled by the library itself (reference: at::zeros_like, rasterization.cpp:156-157).
 *   - `_f32` / `_f64` twins mirror the reference's AT_DISPATCH_FLOATING_TYPES instantiations.
 *   - Scalars keep the reference's C types: `float multiplier, eps, sigmainv` in the kernels
 *     (rasterization.h:31-32).  The `_fused` entry points additionally take `multiplier` and
 *     `boxlen` as double because there they stand for the PYTHON scalars of
 *     `face_vertices_image * multiplier` and `boxlen * multiplier` (dibr.py:32,37-38).
 *   - Return value: DIBR_OK (0) or an error code; dibr_error_string() gives the text.  No C++
 *     exception crosses the boundary.  The Python shim raises RuntimeError on non-zero, which is the
 *     reference's error behaviour (c10::Error -> RuntimeError).
 *   - face_idx semantics are the reference's: int64, index of the face inside its mesh, -1 = none;
 *     ties in depth go to the lowest face index (rasterization_cuda.cu:162-169).
 *
 * Process-wide state (all of it; none affects results)
 *   - two diagnostic counters / hooks: dibr_kernel_launch_count() and dibr_profile_next_launch()
 *     (measurement only, off by default);
 *   - once-per-device cudaFuncSetAttribute calls for the kernels that use > 48 KB of shared memory (made on the
 *     first call per device, so that later calls are CUDA-graph capturable).
 *
 * Tuning switches (kernel experiments; results are identical, see tests/test_gpu_vs_reference.py::
 * test_kernel_variants_give_identical_results).  They are read from the environment ONCE per process and ONLY
 * when DIBR_B200_TUNING=1 is set — without that single switch the library never looks at its environment:
 *     DIBR_RASTER_NT    = 64 | 128 | 256   threads per 16x16 tile of the rasterizer forward kernel (default 64)
 *     DIBR_SOFT_NT      = 64 | 128 | 256   threads per tile of the soft-mask kernels (default 128; 256 for dense
 *                                          scenes: >= 48 faces per 16x16 tile)
 *     DIBR_SOFT_SEL     = n                candidates taken by the first selection pass of a soft-mask tile
 *     DIBR_SOFT_LOG_CAP = n                entries of the soft-mask pair log (default 2 per pixel + 65536); a small
 *                                          value forces the overflow (recompute) path — used by the tests
 */
#ifndef DIBR_B200_H_
#define DIBR_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DIBR_ABI_VERSION 8

enum {
  DIBR_OK = 0,
  DIBR_ERR_INVALID_ARGUMENT = 1, /* null pointer, negative size, unsupported value */
  DIBR_ERR_WORKSPACE = 2,        /* ws == NULL or ws_bytes too small */
  DIBR_ERR_CUDA = 3,             /* a CUDA runtime call / kernel launch failed (dibr_last_cuda_error) */
  DIBR_ERR_TOO_LARGE = 4,        /* a dimension exceeds what the kernels index (H, W <= 32768; B*F < 2^31) */
  DIBR_ERR_SCRATCH_LIMIT = 5     /* a _noalloc entry point needs more than DIBR_NOALLOC_SCRATCH_BYTES of scratch */
};

int dibr_abi_version(void);
const char* dibr_error_string(int code);
/* text of the last CUDA error seen by the calling thread inside this library ("" if none) */
const char* dibr_last_cuda_error(void);

/* Scratch bytes needed by any entry point below for these sizes (a pure function of the sizes).
 * `total_faces` = B*F for the batched layouts, N for the packed one. */
size_t dibr_workspace_bytes(int batch_size, int64_t total_faces, int height, int width);

/* ---------------------------------------------------------------------------------------------
 * rasterize forward, fused layout.
 * Replaces RasterizeCuda.forward end to end (rasterization.py:290-369): valid-face packing,
 * x multiplier, bbox, packed_rasterize_forward_cuda, packed->original index mapping and the -1 fix-up.
 *   z      (B,F,3)   img (B,F,3,2) NDC coords (NOT multiplied)   feat (B,F,3,C)
 *   valid  (B,F) uint8/bool or NULL (= all faces valid)
 *   out_feat (B,H,W,C)   out_face_idx (B,H,W) int64 ORIGINAL face index   out_weights (B,H,W,3) or NULL
 */
int dibr_rasterize_fwd_fused_f32(const float* z, const float* img, const float* feat, const uint8_t* valid,
                                 int B, int F, int C, int H, int W, double multiplier, float eps,
                                 float* out_feat, int64_t* out_face_idx, float* out_weights,
                                 void* ws, size_t ws_bytes, void* stream);
int dibr_rasterize_fwd_fused_f64(const double* z, const double* img, const double* feat, const uint8_t* valid,
                                 int B, int F, int C, int H, int W, double multiplier, float eps,
                                 double* out_feat, int64_t* out_face_idx, double* out_weights,
                                 void* ws, size_t ws_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * packed_rasterize_forward_cuda — reference-shaped (rasterization.h:23-32, rasterization.cpp:49-104).
 *   z (N,3)  img_mult (N,3,2) ALREADY multiplied  bboxes (N,4) [xmin,ymin,xmax,ymax]  feat (N,3,C)
 *   first_idx (B+1) int64 ON DEVICE
 *   out_selected_face_idx (B,H,W): index LOCAL to the mesh's packed range, -1 = none.
 */
int dibr_packed_rasterize_forward_f32(int H, int W, const float* z, const float* img_mult, const float* bboxes,
                                      const float* feat, const int64_t* first_idx, int B, int64_t N, int C,
                                      float multiplier, float eps, float* out_feat,
                                      int64_t* out_selected_face_idx, float* out_weights,
                                      void* ws, size_t ws_bytes, void* stream);
int dibr_packed_rasterize_forward_f64(int H, int W, const double* z, const double* img_mult, const double* bboxes,
                                      const double* feat, const int64_t* first_idx, int B, int64_t N, int C,
                                      float multiplier, float eps, double* out_feat,
                                      int64_t* out_selected_face_idx, double* out_weights,
                                      void* ws, size_t ws_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * rasterize_backward_cuda (rasterization.h:34-41, rasterization.cpp:106-168, kernel
 * rasterization_cuda.cu:238-402).  `interpolated_features` of the reference signature is unused
 * there and is not part of this ABI.
 *   grad_out (B,H,W,C)  face_idx (B,H,W) int64  img (B,F,3,2) NOT multiplied  feat (B,F,3,C)
 *   weights (B,H,W,3), or NULL: the kernel then recomputes them bit-exactly from (pixel, face) with
 *   the forward arithmetic, which needs `multiplier` (ignored when weights != NULL).
 *   grad_img (B,F,3,2), grad_feat (B,F,3,C): zero-filled and accumulated by the library.
 */
int dibr_rasterize_backward_f32(const float* grad_out, const int64_t* face_idx, const float* weights,
                                const float* img, const float* feat, int B, int F, int C, int H, int W,
                                double multiplier, float eps, float* grad_img, float* grad_feat, void* stream);
int dibr_rasterize_backward_f64(const double* grad_out, const int64_t* face_idx, const double* weights,
                                const double* img, const double* feat, int B, int F, int C, int H, int W,
                                double multiplier, float eps, double* grad_img, double* grad_feat, void* stream);

/* ---------------------------------------------------------------------------------------------
 * List-of-features variants.  `rasterize(face_features=[t0, t1, ...])` of the reference concatenates the
 * list (rasterization.py:480-481) and slices the output (:499-505); these entry points consume the n_feat
 * tensors in place and write n_feat separate outputs, so neither the concatenation nor autograd's
 * slice-backward (zero-fill + copy + add per tensor) ever runs.  `feats`, `out_feats`, `grad_outs`,
 * `grad_feats` are HOST arrays of n_feat device pointers, `channels` a HOST array of n_feat ints
 * (n_feat <= 8); tensor t has shape (B,F,3,channels[t]) / (B,H,W,channels[t]).
 */
int dibr_rasterize_fwd_multi_f32(const float* z, const float* img, const float* const* feats, const int* channels,
                                 int n_feat, const uint8_t* valid, int B, int F, int H, int W, double multiplier,
                                 float eps, float* const* out_feats, int64_t* out_face_idx, float* out_weights,
                                 void* ws, size_t ws_bytes, void* stream);
int dibr_rasterize_fwd_multi_f64(const double* z, const double* img, const double* const* feats,
                                 const int* channels, int n_feat, const uint8_t* valid, int B, int F, int H, int W,
                                 double multiplier, float eps, double* const* out_feats, int64_t* out_face_idx,
                                 double* out_weights, void* ws, size_t ws_bytes, void* stream);
int dibr_rasterize_backward_multi_f32(const float* const* grad_outs, const int64_t* face_idx, const float* weights,
                                      const float* img, const float* const* feats, const int* channels, int n_feat,
                                      int B, int F, int H, int W, double multiplier, float eps, float* grad_img,
                                      float* const* grad_feats, void* stream);
int dibr_rasterize_backward_multi_f64(const double* const* grad_outs, const int64_t* face_idx,
                                      const double* weights, const double* img, const double* const* feats,
                                      const int* channels, int n_feat, int B, int F, int H, int W,
                                      double multiplier, float eps, double* grad_img, double* const* grad_feats,
                                      void* stream);

/* ---------------------------------------------------------------------------------------------
 * dibr_soft_mask forward.
 *  _fused   : replaces DibrSoftMaskCuda.forward (dibr.py:29-55): x multiplier, enlarged bboxes, kernel.
 *             img (B,F,3,2) NOT multiplied.
 *  reference-shaped dibr_soft_mask_forward_cuda (dibr_soft_mask.h:23-29): img_mult ALREADY multiplied,
 *             large_bboxes (B,F,4) given.
 *  Outputs: soft (B,H,W); the k-list side buffers prob (B,H,W,K), close_idx (B,H,W,K) int64,
 *  dist_type (B,H,W,K) uint8 are written only when all three are non-NULL (all NULL = skip them;
 *  the backward then recomputes).
 *  keep_pairs != 0 (fused variant, side buffers NULL): a backward will follow — `ws` then has
 *  dibr_soft_mask_workspace_bytes(..., keep_pairs = 1, elem_size) bytes and the kernel also appends every
 *  evaluated (pixel, face, distance type, probability) to a pair log behind the bins (16 bytes per pair, room
 *  for two pairs per pixel; see dibr_soft_mask_bwd_fused, bins_ready == 2).  If the workspace is too small for
 *  the log, or B*H*W >= 2^32, the log is silently not kept and the matching backward recomputes.
 *  fp32 fused variant: the double-precision detour of dibr_soft_mask_cuda.cu:126-138 (`#define EPS 1e-7`) is
 *  taken in fp32 — one approximate reciprocal per edge — and the probability is __expf(-(d * sigmainv / mult^2));
 *  the soft mask stays within 1e-5 relative of the reference-shaped entry point (tested), which keeps the exact
 *  arithmetic because its prob / dist_type outputs are API-visible.
 */
int dibr_soft_mask_fwd_fused_f32(const float* img, const int64_t* face_idx, int B, int F, int H, int W,
                                 float sigmainv, double boxlen, int knum, double multiplier, float* soft,
                                 float* prob, int64_t* close_idx, uint8_t* dist_type,
                                 void* ws, size_t ws_bytes, int keep_pairs, void* stream);
int dibr_soft_mask_fwd_fused_f64(const double* img, const int64_t* face_idx, int B, int F, int H, int W,
                                 float sigmainv, double boxlen, int knum, double multiplier, double* soft,
                                 double* prob, int64_t* close_idx, uint8_t* dist_type,
                                 void* ws, size_t ws_bytes, int keep_pairs, void* stream);
int dibr_soft_mask_forward_f32(const float* img_mult, const float* large_bboxes, const int64_t* face_idx,
                               int B, int F, int H, int W, float sigmainv, int knum, float multiplier,
                               float* soft, float* prob, int64_t* close_idx, uint8_t* dist_type,
                               void* ws, size_t ws_bytes, void* stream);
int dibr_soft_mask_forward_f64(const double* img_mult, const double* large_bboxes, const int64_t* face_idx,
                               int B, int F, int H, int W, float sigmainv, int knum, float multiplier,
                               double* soft, double* prob, int64_t* close_idx, uint8_t* dist_type,
                               void* ws, size_t ws_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * dibr_soft_mask backward.
 *  reference-shaped dibr_soft_mask_backward_cuda (dibr_soft_mask.h:31-40; kernel
 *  dibr_soft_mask_cuda.cu:230-353): consumes the k-list side buffers; img_mult ALREADY multiplied.
 *  _fused (recompute): no side buffers; re-derives each uncovered pixel's first-knum faces, their
 *  probabilities and distance types with the forward arithmetic.  img NOT multiplied.
 *  grad_img (B,F,3,2) is zero-filled and accumulated by the library; it is the gradient w.r.t. the
 *  UN-multiplied coordinates in both variants (the /multiplier is in-kernel, :299-302).
 *  bins_ready != 0 (fused variant): `ws` is the very workspace the matching dibr_soft_mask_fwd_fused call
 *  (same img, sizes, boxlen, multiplier) left behind, untouched since; the setup / binning kernels are then
 *  skipped (the kernels only read the workspace).  0 = rebuild the bins.  bins_ready == 2: that forward call
 *  was made with keep_pairs != 0 and a workspace of dibr_soft_mask_workspace_bytes(..., keep_pairs = 1, ...)
 *  bytes: it logged every evaluated (pixel, face, distance type, probability) behind the bins (12 bytes per
 *  pair, about one pair per pixel), and the backward is then one thread per logged pair instead of a second
 *  derivation of the k-lists; tiles whose pairs did not fit the log are recomputed as with bins_ready == 1.
 */
size_t dibr_soft_mask_workspace_bytes(int batch_size, int64_t total_faces, int height, int width,
                                      int keep_pairs, int elem_size);
int dibr_soft_mask_backward_f32(const float* grad_soft, const float* soft, const int64_t* face_idx,
                                const float* prob, const int64_t* close_idx, const uint8_t* dist_type,
                                const float* img_mult, int B, int F, int H, int W, int knum, float sigmainv,
                                float multiplier, float* grad_img, void* stream);
int dibr_soft_mask_backward_f64(const double* grad_soft, const double* soft, const int64_t* face_idx,
                                const double* prob, const int64_t* close_idx, const uint8_t* dist_type,
                                const double* img_mult, int B, int F, int H, int W, int knum, float sigmainv,
                                float multiplier, double* grad_img, void* stream);
int dibr_soft_mask_bwd_fused_f32(const float* grad_soft, const float* soft, const int64_t* face_idx,
                                 const float* img, int B, int F, int H, int W, float sigmainv, double boxlen,
                                 int knum, double multiplier, float* grad_img,
                                 void* ws, size_t ws_bytes, int bins_ready, void* stream);
int dibr_soft_mask_bwd_fused_f64(const double* grad_soft, const double* soft, const int64_t* face_idx,
                                 const double* img, int B, int F, int H, int W, float sigmainv, double boxlen,
                                 int knum, double multiplier, double* grad_img,
                                 void* ws, size_t ws_bytes, int bins_ready, void* stream);

/* ---------------------------------------------------------------------------------------------
 * dibr_rasterization in ONE call — kaolin/render/mesh/dibr.py:119-209:
 *     rasterize(h, w, z, img, feats, valid_faces = face_normals_z >= 0, multiplier, eps) followed by
 *     dibr_soft_mask(img, face_idx, sigmainv, boxlen, knum, multiplier)
 *  Both face-list sets (the rasterizer's: tight bboxes of the valid faces; the soft mask's: bboxes enlarged by
 *  boxlen * multiplier, all faces) are built from a single read of the coordinates — one memset and three binning
 *  launches instead of two and six — then the two per-tile kernels run.  Results are bit-identical to the two
 *  separate calls (the reference's own requirement, tests/python/kaolin/render/mesh/test_dibr.py:527-529).
 *  `normals_z` (B,F): valid = normals_z >= 0, evaluated in the setup kernel; or `valid` (B,F) uint8 (both NULL = all
 *  faces).  feats / out_feats: HOST arrays of n_feat device pointers, channels: HOST array (as in _multi).
 *  ws: dibr_rasterization_workspace_bytes(B, B*F, H, W, keep_pairs) bytes.  keep_pairs != 0: the soft-mask kernel also
 *  logs its evaluated pairs into `ws` for the backward call below (which must then get the same `ws`, untouched).
 *  _bwd: grad_outs (HOST array; NULL or all-NULL = no upstream feature gradient) and grad_soft (NULL = none) ->
 *  grad_img (B,F,3,2) = the SUM of both paths' gradients, accumulated in place by both kernels (zero-filled once by
 *  the library), and grad_feats.  bins_ready: 6 = `ws` comes from the forward call with keep_pairs (bins + log);
 *  5 = bins only; 0 = `ws` is scratch of dibr_soft_mask_workspace_bytes(..., 0, ...) bytes, bin again.
 */
size_t dibr_rasterization_workspace_bytes(int batch_size, int64_t total_faces, int height, int width, int keep_pairs);
int dibr_rasterization_fwd_fused_f32(const float* z, const float* img, const float* const* feats,
                                     const int* channels, int n_feat, const float* normals_z, const uint8_t* valid,
                                     int B, int F, int H, int W, double multiplier, float eps, float sigmainv,
                                     double boxlen, int knum, float* const* out_feats, int64_t* out_face_idx,
                                     float* soft, void* ws, size_t ws_bytes, int keep_pairs, void* stream);
int dibr_rasterization_fwd_fused_f64(const double* z, const double* img, const double* const* feats,
                                     const int* channels, int n_feat, const double* normals_z, const uint8_t* valid,
                                     int B, int F, int H, int W, double multiplier, float eps, float sigmainv,
                                     double boxlen, int knum, double* const* out_feats, int64_t* out_face_idx,
                                     double* soft, void* ws, size_t ws_bytes, int keep_pairs, void* stream);
int dibr_rasterization_bwd_fused_f32(const float* const* grad_outs, const float* grad_soft, const float* soft,
                                     const int64_t* face_idx, const float* img, const float* const* feats,
                                     const int* channels, int n_feat, int B, int F, int H, int W, double multiplier,
                                     float eps, float sigmainv, double boxlen, int knum, float* grad_img,
                                     float* const* grad_feats, void* ws, size_t ws_bytes, int bins_ready,
                                     void* stream);
int dibr_rasterization_bwd_fused_f64(const double* const* grad_outs, const double* grad_soft, const double* soft,
                                     const int64_t* face_idx, const double* img, const double* const* feats,
                                     const int* channels, int n_feat, int B, int F, int H, int W, double multiplier,
                                     float eps, float sigmainv, double boxlen, int knum, double* grad_img,
                                     double* const* grad_feats, void* ws, size_t ws_bytes, int bins_ready,
                                     void* stream);

/* ---------------------------------------------------------------------------------------------
 * dibr_rasterization straight from the mesh — prepare_vertices folded into K0:
 *     kaolin/render/mesh/utils.py:128-175 (prepare_vertices: rotate_translate_points / padded @ camera_transform,
 *     perspective_camera, index_vertices_by_faces, face_normals) followed by kaolin/render/mesh/dibr.py:119-209
 *     (dibr_rasterization with face_normals_z = face_normals[:, :, 2]).
 *  The setup kernel gathers each face's vertices through `faces` (F,3) int64 from `vertices` (Bv,V,3) (Bv = 1: one
 *  mesh seen by B cameras; Bv = B: one mesh per camera), applies the camera (rot (B,3,3) + trans (B,3), or transform
 *  (B,4,3); exactly one form) and camera_proj (3), culls with the unit normal's z and bins — the camera-space
 *  (B,F,3,3) tensor and the (B,F,3) normals are never materialised.  It WRITES face_vertices_image (B,F,3,2) and
 *  face_vertices_z (B,F,3) (caller's buffers): the per-tile kernels and the backward read them, and they are
 *  bit-identical to dibr_prepare_vertices_fwd's outputs, so every result equals the unfused chain's.
 *  ws / keep_pairs / feats / out_feats: as dibr_rasterization_fwd_fused.
 *  _bwd: the image-plane gradients of both outputs are accumulated in grad_image_scratch (B,F,3,2; caller's scratch,
 *  contents unspecified afterwards) and folded by prepare_vertices' backward into grad_vertices (Bv,V,3) and, when
 *  non-NULL, grad_proj (3) / grad_rot (B,3,3) / grad_trans (B,3) / grad_transform (B,4,3) — all zero-filled by the
 *  library; at least one of them must be given.  face_vertices_z gets no gradient (rasterization_cuda.cu:238-402
 *  has none either).  bins_ready as dibr_rasterization_bwd_fused.
 */
int dibr_rasterization_from_mesh_fwd_f32(const float* vertices, const int64_t* faces, const float* camera_proj,
                                         const float* rot, const float* trans, const float* transform, int Bv, int V,
                                         const float* const* feats, const int* channels, int n_feat, int B, int F,
                                         int H, int W, double multiplier, float eps, float sigmainv, double boxlen,
                                         int knum, float* face_vertices_image, float* face_vertices_z,
                                         float* const* out_feats, int64_t* out_face_idx, float* soft, void* ws,
                                         size_t ws_bytes, int keep_pairs, void* stream);
int dibr_rasterization_from_mesh_fwd_f64(const double* vertices, const int64_t* faces, const double* camera_proj,
                                         const double* rot, const double* trans, const double* transform, int Bv,
                                         int V, const double* const* feats, const int* channels, int n_feat, int B,
                                         int F, int H, int W, double multiplier, float eps, float sigmainv,
                                         double boxlen, int knum, double* face_vertices_image,
                                         double* face_vertices_z, double* const* out_feats, int64_t* out_face_idx,
                                         double* soft, void* ws, size_t ws_bytes, int keep_pairs, void* stream);
int dibr_rasterization_from_mesh_bwd_f32(const float* const* grad_outs, const float* grad_soft, const float* soft,
                                         const int64_t* face_idx, const float* face_vertices_image,
                                         const float* const* feats, const int* channels, int n_feat,
                                         const float* vertices, const int64_t* faces, const float* camera_proj,
                                         const float* rot, const float* trans, const float* transform, int Bv, int V,
                                         int B, int F, int H, int W, double multiplier, float eps, float sigmainv,
                                         double boxlen, int knum, float* grad_image_scratch,
                                         float* const* grad_feats, float* grad_vertices, float* grad_proj,
                                         float* grad_rot, float* grad_trans, float* grad_transform, void* ws,
                                         size_t ws_bytes, int bins_ready, void* stream);
int dibr_rasterization_from_mesh_bwd_f64(const double* const* grad_outs, const double* grad_soft, const double* soft,
                                         const int64_t* face_idx, const double* face_vertices_image,
                                         const double* const* feats, const int* channels, int n_feat,
                                         const double* vertices, const int64_t* faces, const double* camera_proj,
                                         const double* rot, const double* trans, const double* transform, int Bv,
                                         int V, int B, int F, int H, int W, double multiplier, float eps,
                                         float sigmainv, double boxlen, int knum, double* grad_image_scratch,
                                         double* const* grad_feats, double* grad_vertices, double* grad_proj,
                                         double* grad_rot, double* grad_trans, double* grad_transform, void* ws,
                                         size_t ws_bytes, int bins_ready, void* stream);

/* ---------------------------------------------------------------------------------------------
 * prepare_vertices — the producer of the path's inputs (first "next" row).
 *  Replaces the torch composition of kaolin/render/mesh/utils.py:128-175:
 *    camera transform      rotate_translate_points (kaolin/render/camera/legacy.py:22-37) when rot/trans are
 *                          given, or [v,1] @ camera_transform (B,4,3) (utils.py:163-167) when transform is given
 *    perspective_camera    kaolin/render/camera/legacy.py:120-139   (x*p0/(z*p2), y*p1/(z*p2))
 *    index_vertices_by_faces (twice)   kaolin/ops/mesh/mesh.py:24-45
 *    face_normals(unit=True)           kaolin/ops/mesh/trianglemesh.py:313-336  (n / (|n| + 1e-10))
 *  by one kernel (thread = view x face).  vertices (Bv,V,3) with Bv == 1 (one mesh broadcast over the
 *  B cameras, as torch broadcasting does in the reference) or Bv == B; faces (F,3) int64 in [0,V)
 *  (an out-of-range index yields NaN outputs for that face, never an out-of-bounds read);
 *  camera_proj (3); exactly one of {rot (B,3,3) + trans (B,3)} or {transform (B,4,3)} non-NULL.
 *  Outputs: face_vertices_camera (B,F,3,3), face_vertices_image (B,F,3,2), face_normals (B,F,3).
 *  _bwd: any of the three upstream gradients may be NULL (= zeros); grad_vertices (Bv,V,3) is zero-filled
 *  and accumulated by the library (summed over views when Bv == 1).  grad_camera_proj (3), grad_rot (B,3,3),
 *  grad_trans (B,3), grad_transform (B,4,3): optional gradients of the camera tensors (NULL = not wanted; only
 *  those of the camera form in use may be non-NULL; zero-filled and accumulated by the library) — the reference
 *  is pure torch and differentiates through all of them.  grad_vertices may be NULL when a camera gradient is asked.
 *  Floating point: agrees with the reference within 1e-5 relative (fp32) — its matmul has no fixed order.
 */
int dibr_prepare_vertices_fwd_f32(const float* vertices, const int64_t* faces, const float* camera_proj,
                                  const float* rot, const float* trans, const float* transform,
                                  int B, int Bv, int V, int F, float* face_vertices_camera,
                                  float* face_vertices_image, float* face_normals, void* stream);
int dibr_prepare_vertices_fwd_f64(const double* vertices, const int64_t* faces, const double* camera_proj,
                                  const double* rot, const double* trans, const double* transform,
                                  int B, int Bv, int V, int F, double* face_vertices_camera,
                                  double* face_vertices_image, double* face_normals, void* stream);
int dibr_prepare_vertices_bwd_f32(const float* grad_camera, const float* grad_image, const float* grad_normals,
                                  const float* vertices, const int64_t* faces, const float* camera_proj,
                                  const float* rot, const float* trans, const float* transform,
                                  int B, int Bv, int V, int F, float* grad_vertices, float* grad_camera_proj,
                                  float* grad_rot, float* grad_trans, float* grad_transform, void* stream);
int dibr_prepare_vertices_bwd_f64(const double* grad_camera, const double* grad_image,
                                  const double* grad_normals, const double* vertices, const int64_t* faces,
                                  const double* camera_proj, const double* rot, const double* trans,
                                  const double* transform, int B, int Bv, int V, int F,
                                  double* grad_vertices, double* grad_camera_proj, double* grad_rot,
                                  double* grad_trans, double* grad_transform, void* stream);

/* ---------------------------------------------------------------------------------------------
 * texture_mapping — the consumer of the rasterizer's interpolated uv map.
 *  Replaces kaolin/render/mesh/utils.py:23-76 (clamp to [0,1], *2-1, flip y,
 *  torch.nn.functional.grid_sample(mode, align_corners=False, padding_mode='border'), permute to
 *  channels-last) and, when `mask` is non-NULL, the tutorial epilogue torch.clamp(image * mask, 0., 1.)
 *  (examples/tutorial/ian_dibr.py:254-258) by one kernel per direction.
 *  uv (B,N,2) with N = h*w (dense) or num_points (sparse); texture_maps (Bt,C,Ht,Wt) with Bt == B, or
 *  Bt == 1 for one texture shared by the batch; mask (B,N) or NULL; mode 0 = nearest, 1 = bilinear;
 *  out (B,N,C).  _bwd: grad_uv (B,N,2), grad_texture_maps (Bt,C,Ht,Wt; zero-filled and accumulated by the
 *  library), grad_mask (B,N) — each may be NULL when not needed (grad_mask requires mask).
 */
int dibr_texture_mapping_fwd_f32(const float* uv, const float* texture_maps, const float* mask, int B,
                                 int64_t N, int Bt, int C, int Ht, int Wt, int mode, float* out, void* stream);
int dibr_texture_mapping_fwd_f64(const double* uv, const double* texture_maps, const double* mask, int B,
                                 int64_t N, int Bt, int C, int Ht, int Wt, int mode, double* out, void* stream);
int dibr_texture_mapping_bwd_f32(const float* grad_out, const float* uv, const float* texture_maps,
                                 const float* mask, int B, int64_t N, int Bt, int C, int Ht, int Wt, int mode,
                                 float* grad_uv, float* grad_texture_maps, float* grad_mask, void* stream);
int dibr_texture_mapping_bwd_f64(const double* grad_out, const double* uv, const double* texture_maps,
                                 const double* mask, int B, int64_t N, int Bt, int C, int Ht, int Wt, int mode,
                                 double* grad_uv, double* grad_texture_maps, double* grad_mask, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Adam step of the fitting loop — torch.optim.Adam(params, lr, betas, eps) without weight decay / amsgrad
 * (examples/tutorial/ian_dibr.py:116-123, `vertices_optim.step()` at :287) for ONE parameter tensor of n elements,
 * one thread per element (torch hands a tensor to one CTA per 65 536 elements: a 20 k-vertex mesh is a single CTA).
 * In place on param / exp_avg / exp_avg_sq.  `step` (1 float) is the DEVICE step counter, read and then incremented
 * by the kernel (graph-capturable: nothing on the host changes between steps); `arrivals` (1 unsigned) is scratch that
 * the caller zero-initialises once.  Same update as torch: m = b1 m + (1 - b1) g; v = b2 v + (1 - b2) g^2;
 * param -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps), t = step + 1, arithmetic in double.
 */
int dibr_adam_step_f32(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, float* step,
                       unsigned* arrivals, int64_t n, double lr, double beta1, double beta2, double eps,
                       void* stream);
int dibr_adam_step_f64(double* param, const double* grad, double* exp_avg, double* exp_avg_sq, float* step,
                       unsigned* arrivals, int64_t n, double lr, double beta1, double beta2, double eps,
                       void* stream);

/* ---------------------------------------------------------------------------------------------
 * Image L1 loss of the fitting loop — examples/tutorial/ian_dibr.py:260  image_loss = torch.mean(torch.abs(image -
 * gt_image)): loss (device scalar) = mean over the n elements of |a - b|, as a deterministic two-stage reduction (no
 * float atomics); ws: dibr_image_l1_workspace_bytes(n, elem_size) bytes.  _bwd: grad_loss is a DEVICE scalar;
 * grad_a = sign(a - b) * grad_loss / n, grad_b = -grad_a (either may be NULL).
 */
size_t dibr_image_l1_workspace_bytes(int64_t n, int elem_size);
int dibr_image_l1_fwd_f32(const float* a, const float* b, int64_t n, float* loss, void* ws, size_t ws_bytes,
                          void* stream);
int dibr_image_l1_fwd_f64(const double* a, const double* b, int64_t n, double* loss, void* ws, size_t ws_bytes,
                          void* stream);
int dibr_image_l1_bwd_f32(const float* grad_loss, const float* a, const float* b, int64_t n, float* grad_a,
                          float* grad_b, void* stream);
int dibr_image_l1_bwd_f64(const double* grad_loss, const double* a, const double* b, int64_t n, double* grad_a,
                          double* grad_b, void* stream);

/* ---------------------------------------------------------------------------------------------
 * mask_iou — the consumer of soft_mask.
 *  Replaces kaolin/metrics/render.py:18-40: loss = 1 - mean_b( sum(l*r) / (sum(l + r - l*r) + 1e-10) ).
 *  lhs, rhs (B,P) with P = height*width.  _fwd writes the scalar loss (device) and sums (B,2) = per-view
 *  {sum(l*r), sum(l+r-l*r)}, which _bwd consumes; the reduction is two-stage and deterministic (no float
 *  atomics).  ws: dibr_mask_iou_workspace_bytes(B, P, elem_size) bytes.  _bwd: grad_loss is a DEVICE scalar;
 *  grad_lhs / grad_rhs (B,P) may each be NULL.
 */
size_t dibr_mask_iou_workspace_bytes(int B, int64_t P, int elem_size);
int dibr_mask_iou_fwd_f32(const float* lhs, const float* rhs, int B, int64_t P, float* loss, float* sums,
                          void* ws, size_t ws_bytes, void* stream);
int dibr_mask_iou_fwd_f64(const double* lhs, const double* rhs, int B, int64_t P, double* loss, double* sums,
                          void* ws, size_t ws_bytes, void* stream);
/* mask_iou forward WITHOUT a caller workspace and without any allocation.
 *  No counterpart in the reference: added for callers that cannot call
 *  cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs and
 *  results as dibr_mask_iou_fwd_* (it runs the same two kernels); the scratch of the two-stage reduction lives in a
 *  file-scope `static __device__` array of DIBR_NOALLOC_SCRATCH_BYTES inside the library, which every GPU that loads
 *  the library holds until the process exits.  The call makes no cudaMalloc / cudaFree / allocating library call; the
 *  first call per device takes the array's device address with cudaGetSymbolAddress (set-up), later calls only launch.
 *  Limit: dibr_mask_iou_workspace_bytes(B, P, elem_size) <= dibr_noalloc_scratch_bytes(), i.e. B * ceil(P / 8192)
 *  <= 131072 (fp32) / 65536 (fp64) — e.g. 128 views of 2048 x 2048 in fp32; beyond it the call returns
 *  DIBR_ERR_SCRATCH_LIMIT before touching the GPU (use dibr_mask_iou_fwd_* with a workspace instead).
 *  The array is shared by every call on the device and keeps the last call's contents; nothing in it must start at
 *  zero (each call writes every slot it reads).  Calls that OVERLAP on the GPU — different streams without an
 *  ordering between them — race on it: serialise them.  On one stream, back-to-back calls are safe.
 */
#define DIBR_NOALLOC_SCRATCH_BYTES (1u << 20)
size_t dibr_noalloc_scratch_bytes(void);
int dibr_mask_iou_fwd_noalloc_f32(const float* lhs, const float* rhs, int B, int64_t P, float* loss, float* sums,
                                  void* stream);
int dibr_mask_iou_fwd_noalloc_f64(const double* lhs, const double* rhs, int B, int64_t P, double* loss,
                                  double* sums, void* stream);
int dibr_mask_iou_bwd_f32(const float* grad_loss, const float* lhs, const float* rhs, const float* sums,
                          int B, int64_t P, float* grad_lhs, float* grad_rhs, void* stream);
int dibr_mask_iou_bwd_f64(const double* grad_loss, const double* lhs, const double* rhs, const double* sums,
                          int B, int64_t P, double* grad_lhs, double* grad_rhs, void* stream);

/* ---------------------------------------------------------------------------------------------
 * deftet_sparse_render — DefTet volumetric renderer on the path's bins.
 *  For every query point: the first knum faces IN FACE ORDER that contain the point (half-open bbox test,
 *  barycentric weights >= 0) with depth in [render_ranges[...,0], render_ranges[...,1]).
 *  reference-shaped dibr_deftet_forward_* = deftet_sparse_render_forward_cuda (kaolin/csrc/render/mesh/deftet.h,
 *    deftet.cpp:47-109; kernel deftet_cuda.cu:31-190): face_bboxes (B,F,4) given; outputs in face order with the
 *    reference's fills: face_idx (B,P,knum) int64 (-1), pixel_depth (-inf), w0, w1 (0).
 *  _fused = DeftetSparseRenderer.forward (kaolin/render/mesh/deftet.py:272-318): bboxes computed in-kernel,
 *    entries sorted by depth (descending; equal depths keep face order), features of n_feat tensors
 *    (B,F,3,channels[t]) interpolated into outs[t] (B,P,knum,channels[t]) (no torch.cat), weights (B,P,knum,3)
 *    optional (backward needs them).
 *  dibr_deftet_backward_* = deftet_sparse_render_backward_cuda (deftet.cpp:111-170, deftet_cuda.cu:240-410) for
 *    a list of feature tensors; grad_img (B,F,3,2) and grad_feats[t] are zero-filled and accumulated by the
 *    library; grad_img or any grad_feats[t] may be NULL.
 *  knum <= 4064.  ws: dibr_deftet_workspace_bytes(B, total_faces).  eps: float, as in the reference.
 */
size_t dibr_deftet_workspace_bytes(int batch_size, int64_t total_faces);
int dibr_deftet_forward_f32(const float* face_vertices_z, const float* face_vertices_image, const float* face_bboxes,
                            const float* pixel_coords, const float* render_ranges, int B, int F, int P, int knum,
                            float eps, int64_t* face_idx, float* pixel_depth, float* w0, float* w1,
                            void* ws, size_t ws_bytes, void* stream);
int dibr_deftet_forward_f64(const double* face_vertices_z, const double* face_vertices_image,
                            const double* face_bboxes, const double* pixel_coords, const double* render_ranges,
                            int B, int F, int P, int knum, float eps, int64_t* face_idx, double* pixel_depth,
                            double* w0, double* w1, void* ws, size_t ws_bytes, void* stream);
int dibr_deftet_render_fused_f32(const float* pixel_coords, const float* render_ranges, const float* face_vertices_z,
                                 const float* face_vertices_image, const float* const* feats, const int* channels,
                                 int n_feat, int B, int F, int P, int knum, float eps, float* const* outs,
                                 int64_t* face_idx, float* weights, void* ws, size_t ws_bytes, void* stream);
int dibr_deftet_render_fused_f64(const double* pixel_coords, const double* render_ranges,
                                 const double* face_vertices_z, const double* face_vertices_image,
                                 const double* const* feats, const int* channels, int n_feat, int B, int F, int P,
                                 int knum, float eps, double* const* outs, int64_t* face_idx, double* weights,
                                 void* ws, size_t ws_bytes, void* stream);
int dibr_deftet_backward_f32(const float* const* grad_outs, const int64_t* face_idx, const float* weights,
                             const float* face_vertices_image, const float* const* feats, const int* channels,
                             int n_feat, int B, int F, int P, int knum, float eps, float* grad_img,
                             float* const* grad_feats, void* stream);
int dibr_deftet_backward_f64(const double* const* grad_outs, const int64_t* face_idx, const double* weights,
                             const double* face_vertices_image, const double* const* feats, const int* channels,
                             int n_feat, int B, int F, int P, int knum, float eps, double* grad_img,
                             double* const* grad_feats, void* stream);

/* Counters for bench.py's `gpu_launches`: number of kernels this library has launched from the
 * calling process since load / since the last reset (relaxed atomic; diagnostics only). */
uint64_t dibr_kernel_launch_count(void);
void dibr_reset_kernel_launch_count(void);

/* Measurement hook (bench.py roofline leg).  Brackets the calling thread's NEXT launch of kernel
 * `kernel_id` with cudaEventRecord(start_event) / cudaEventRecord(stop_event) on the launch stream, then
 * disarms itself.  kernel_id < 0 disarms.  The events are cudaEvent_t handles owned by the caller. */
enum {
  DIBR_KERNEL_SETUP_COUNT = 0,
  DIBR_KERNEL_SCAN = 1,
  DIBR_KERNEL_FILL = 2,
  DIBR_KERNEL_RASTER_FWD = 3,
  DIBR_KERNEL_RASTER_BWD = 4,
  DIBR_KERNEL_SOFT_FWD = 5,
  DIBR_KERNEL_SOFT_BWD = 6,
  DIBR_KERNEL_PREPARE_FWD = 7,
  DIBR_KERNEL_PREPARE_BWD = 8,
  DIBR_KERNEL_TEXTURE_FWD = 9,
  DIBR_KERNEL_TEXTURE_BWD = 10,
  DIBR_KERNEL_IOU_FWD = 11,
  DIBR_KERNEL_IOU_BWD = 12,
  DIBR_KERNEL_DEFTET_FWD = 13,
  DIBR_KERNEL_DEFTET_BWD = 14,
  DIBR_KERNEL_L1_FWD = 15,
  DIBR_KERNEL_L1_BWD = 16,
  DIBR_KERNEL_COUNT = 17
};
int dibr_profile_next_launch(int kernel_id, void* start_event, void* stop_event);

#ifdef __cplusplus
}
#endif
#endif /* DIBR_B200_H_ */
