// dibr_c_abi.cu — extern "C" entry points of libdibr_b200.so (declared in include/dibr_b200.h).
// Host logic only: argument checks, workspace carving, kernel launches on the caller's stream.
#include "../../include/dibr_b200.h"
#include "dibr_binning.cuh"
#include "dibr_raster_fwd.cuh"
#include "dibr_raster_bwd.cuh"
#include "dibr_soft_mask.cuh"
#include "dibr_prepare.cuh"
#include "dibr_texture.cuh"
#include "dibr_losses.cuh"
#include "dibr_deftet.cuh"

#include <nvtx3/nvToolsExt.h>       // header-only NVTX v3: ranges cost a load + branch unless a tool (nsys, ncu) is attached

#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>

using namespace dibr;

namespace {

std::atomic<uint64_t> g_launches{0};
thread_local char t_cuda_err[256] = {0};

inline int cuda_fail(cudaError_t e, const char* where) {
  snprintf(t_cuda_err, sizeof(t_cuda_err), "%s: %s", where, cudaGetErrorString(e));
  return DIBR_ERR_CUDA;
}
#define CUDA_TRY(expr)                                  \
  do {                                                  \
    cudaError_t _e = (expr);                            \
    if (_e != cudaSuccess) return cuda_fail(_e, #expr); \
  } while (0)
#define LAUNCH_CHECK(name)                                  \
  do {                                                      \
    g_launches.fetch_add(1, std::memory_order_relaxed);     \
    cudaError_t _e = cudaGetLastError();                    \
    if (_e != cudaSuccess) return cuda_fail(_e, name);      \
  } while (0)

// Optional per-kernel timing hook (bench.py's roofline leg): the caller names ONE kernel id and a
// pair of CUDA events; the next launch of that kernel (from any thread) is bracketed by the two events on
// the launch stream.  Off by default; costs nothing when unset.
// Process-wide (forward runs on the Python thread, backward on the autograd thread); measurement only.
std::atomic<int> g_prof_kernel{-1};
std::atomic<cudaEvent_t> g_prof_start{nullptr}, g_prof_stop{nullptr};
const char* const kKernelNames[DIBR_KERNEL_COUNT] = {
    "dibr/setup_count", "dibr/scan", "dibr/fill", "dibr/raster_fwd", "dibr/raster_bwd", "dibr/soft_mask_fwd",
    "dibr/soft_mask_bwd", "dibr/prepare_vertices_fwd", "dibr/prepare_vertices_bwd", "dibr/texture_mapping_fwd",
    "dibr/texture_mapping_bwd", "dibr/mask_iou_fwd", "dibr/mask_iou_bwd", "dibr/deftet_fwd", "dibr/deftet_bwd",
    "dibr/image_l1_fwd", "dibr/image_l1_bwd"};
// Every kernel launch site sits in one of these scopes: an NVTX range named after the kernel (so that an nsys / ncu
// timeline of a training step reads "dibr/raster_fwd ..." instead of mangled template names) + the timing hook.
struct ProfScope {
  bool on, nvtx; cudaStream_t s;
  ProfScope(int id, cudaStream_t st) : on(false), nvtx(id >= 0 && id < DIBR_KERNEL_COUNT), s(st) {
    if (nvtx) nvtxRangePushA(kKernelNames[id]);
    int expect = id;
    if (g_prof_kernel.load(std::memory_order_acquire) == id &&
        g_prof_kernel.compare_exchange_strong(expect, -2)) {          // claim it (one launch only)
      on = true;
      cudaEventRecord(g_prof_start.load(), s);
    }
  }
  ~ProfScope() {
    if (on) { cudaEventRecord(g_prof_stop.load(), s); g_prof_kernel.store(-1, std::memory_order_release); }
    if (nvtx) nvtxRangePop();
  }
};

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

// Kernel-tuning knobs (include/dibr_b200.h, "Tuning switches"): read once per process, and only when the single
// documented switch DIBR_B200_TUNING=1 is set — a plain process never changes behaviour through its environment.
inline const char* tuning_env(const char* name) {
  static const bool on = [] { const char* e = getenv("DIBR_B200_TUNING"); return e && e[0] == '1' && e[1] == 0; }();
  return on ? getenv(name) : nullptr;
}

// Workspace carving.  One BinLists = a ZERO part (counters: one memset) and a REST part (offsets, order, ranges, lists).
// Standalone calls use [zero | rest].  The dibr_rasterization call (and every soft-mask call, with an empty rasterizer
// part) uses the FUSED layout [soft zero | raster zero | log zero | soft rest | raster rest | log rest]: all
// counters of both list sets and of the pair log are contiguous and are cleared by ONE memset.
struct WsLayout {
  size_t ranges, tile_count, tile_offset, tile_cursor, pairs, large_list, large_count, alloc_cursor, tile_order;
  size_t bucket_hist, bucket_cursor;
  size_t zero_begin, zero_end;     // [tile_count | large_count | alloc_cursor | bucket_hist | bucket_cursor]
  size_t total;
};
size_t ws_zero_part(WsLayout& w, int B, const Geometry& geo, size_t o) {
  const size_t bt = (size_t)B * geo.slots_per_view;
  w.zero_begin = o;
  w.tile_count = o;  o = align_up(o + bt * sizeof(int));
  w.large_count = o; o = align_up(o + (size_t)B * sizeof(int));
  w.alloc_cursor = o; o = align_up(o + sizeof(int));
  w.bucket_hist = o; o = align_up(o + (size_t)B * NBUCKET * sizeof(int));
  w.bucket_cursor = o; o = align_up(o + (size_t)B * NBUCKET * sizeof(int));
  w.zero_end = o;
  return o;
}
size_t ws_rest_part(WsLayout& w, int B, int64_t Ntot, const Geometry& geo, size_t o) {
  const size_t bt = (size_t)B * geo.slots_per_view;
  w.tile_offset = o; o = align_up(o + bt * sizeof(int));
  w.tile_cursor = o; o = align_up(o + bt * sizeof(int));
  w.tile_order = o;  o = align_up(o + (size_t)B * geo.tiles_per_view * sizeof(int));
  w.ranges = o;      o = align_up(o + (size_t)Ntot * sizeof(uint2));
  w.large_list = o;  o = align_up(o + (size_t)Ntot * sizeof(int));
  w.pairs = o;       o = align_up(o + (size_t)Ntot * PAIRS_PER_FACE * sizeof(int));
  w.total = o;
  return o;
}
WsLayout ws_layout(int B, int64_t Ntot, const Geometry& geo) {
  WsLayout w;
  ws_rest_part(w, B, Ntot, geo, ws_zero_part(w, B, geo, 0));
  return w;
}
BinLists carve(void* ws, const WsLayout& w) {
  unsigned char* p = static_cast<unsigned char*>(ws);
  BinLists L;
  L.ranges = reinterpret_cast<uint2*>(p + w.ranges);
  L.tile_count = reinterpret_cast<int*>(p + w.tile_count);
  L.tile_offset = reinterpret_cast<int*>(p + w.tile_offset);
  L.tile_cursor = reinterpret_cast<int*>(p + w.tile_cursor);
  L.pairs = reinterpret_cast<int*>(p + w.pairs);
  L.large_list = reinterpret_cast<int*>(p + w.large_list);
  L.large_count = reinterpret_cast<int*>(p + w.large_count);
  L.alloc_cursor = reinterpret_cast<int*>(p + w.alloc_cursor);
  L.tile_order = reinterpret_cast<int*>(p + w.tile_order);
  L.bucket_hist = reinterpret_cast<int*>(p + w.bucket_hist);
  L.bucket_cursor = reinterpret_cast<int*>(p + w.bucket_cursor);
  return L;
}

int check_dims(int B, int64_t Ntot, int H, int W) {
  if (B < 0 || Ntot < 0 || H < 0 || W < 0) return DIBR_ERR_INVALID_ARGUMENT;
  if (H > 32768 || W > 32768) return DIBR_ERR_TOO_LARGE;
  if (Ntot >= (int64_t)1 << 31 || (int64_t)B * H * W >= (int64_t)1 << 40) return DIBR_ERR_TOO_LARGE;
  const Geometry geo = make_geometry(B, H, W);
  if ((int64_t)B * geo.slots_per_view >= (int64_t)1 << 31) return DIBR_ERR_TOO_LARGE;
  return DIBR_OK;
}

// K0 + K1 on one set of lists: ranges, counts, scan, fill (+ tile order).  3 kernels + 1 memset.
template <typename T>
int run_binning(const SetupParams<T>& sp, const FaceLayout& fl, const Geometry& geo, void* ws, size_t ws_bytes,
                cudaStream_t s, BinLists& L) {
  const WsLayout w = ws_layout(geo.B, fl.Ntot, geo);
  if (!ws || ws_bytes < w.total) return DIBR_ERR_WORKSPACE;
  L = carve(ws, w);
  CUDA_TRY(cudaMemsetAsync(static_cast<unsigned char*>(ws) + w.zero_begin, 0, w.zero_end - w.zero_begin, s));
  BinLists2 LL; LL.l[0] = L; LL.l[1] = L;
  const unsigned nb = binning_blocks(fl, geo, false);
  if (nb > 0) {
    { ProfScope ps(DIBR_KERNEL_SETUP_COUNT, s); setup_count_kernel<T, false, false><<<nb, 256, 0, s>>>(sp, (T)0, fl, geo, LL, PrepareParams<T>(), MeshOut<T>()); }
    LAUNCH_CHECK("setup_count_kernel");
  }
  { ProfScope ps(DIBR_KERNEL_SCAN, s); scan_kernel<<<scan_grid(geo, 1), SCAN_CHUNK, 0, s>>>(geo, LL); }
  LAUNCH_CHECK("scan_kernel");
  { ProfScope ps(DIBR_KERNEL_FILL, s); fill_kernel<false><<<binning_blocks(fl, geo, true), 256, 0, s>>>(fl, geo, LL); }
  LAUNCH_CHECK("fill_kernel");
  return DIBR_OK;
}

template <typename T>
int make_featsets(const T* const* in, T* const* out, const int* channels, int n, bool need_in,
                  FeatSet<const T>& fi, FeatSet<T>& fo) {
  if (n < 0 || n > MAX_FEATS || (n > 0 && (!channels || !out || !in))) return DIBR_ERR_INVALID_ARGUMENT;
  fi.n = fo.n = n;
  for (int t = 0; t < MAX_FEATS; ++t) { fi.ptr[t] = nullptr; fo.ptr[t] = nullptr; fi.ch[t] = fo.ch[t] = 0; }
  for (int t = 0; t < n; ++t) {
    if (channels[t] < 0) return DIBR_ERR_INVALID_ARGUMENT;
    if (channels[t] > 0 && (!out[t] || (need_in && !in[t]))) return DIBR_ERR_INVALID_ARGUMENT;
    fi.ptr[t] = in[t]; fo.ptr[t] = out[t]; fi.ch[t] = fo.ch[t] = channels[t];
  }
  return DIBR_OK;
}

template <typename T>
int rasterize_fwd_impl(const T* z, const T* img, const T* bbox_in, const T* const* feats, const int* channels,
                       int n_feat, const uint8_t* valid, const int64_t* first_idx, int B, int F, int64_t Ntot,
                       int H, int W, T coord_mult, int premultiplied, float pix_mult, float eps,
                       T* const* out_feats, int64_t* out_idx, T* out_w, void* ws, size_t ws_bytes, void* stream) {
  if (int e = check_dims(B, Ntot, H, W)) return e;
  if (!(pix_mult > 0.0f)) return DIBR_ERR_INVALID_ARGUMENT;
  if (B == 0 || H == 0 || W == 0) return DIBR_OK;
  if ((Ntot > 0 && (!z || !img)) || !out_idx) return DIBR_ERR_INVALID_ARGUMENT;
  FeatSet<const T> fin; FeatSet<T> fout;
  if (int e = make_featsets<T>(feats, out_feats, channels, n_feat, Ntot > 0, fin, fout)) return e;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const Geometry geo = make_geometry(B, H, W);
  FaceLayout fl; fl.first_idx = first_idx; fl.F = F; fl.Ntot = Ntot; fl.B = B;
  SetupParams<T> sp;
  sp.img = img; sp.bbox_in = bbox_in; sp.valid = valid; sp.normals_z = nullptr; sp.coord_mult = coord_mult;
  sp.premultiplied = premultiplied; sp.margin = (T)0; sp.pix_mult = pix_mult;
  BinLists L;
  if (int e = run_binning<T>(sp, fl, geo, ws, ws_bytes, s, L)) return e;
  RasterFwdParams<T> rp;
  rp.z = z; rp.img = img; rp.feat = fin; rp.coord_mult = coord_mult; rp.premultiplied = premultiplied;
  rp.pix_mult = pix_mult; rp.eps = eps; rp.out_feat = fout; rp.out_idx = out_idx; rp.out_w = out_w;
  rp.soft_out = nullptr; rp.tile_unc = nullptr;
  {
    ProfScope ps(DIBR_KERNEL_RASTER_FWD, s);
    launch_raster_fwd(rp, fl, geo, L, s);
  }
  LAUNCH_CHECK("raster_fwd_kernel");
  return DIBR_OK;
}

template <typename T>
int rasterize_bwd_impl(const T* const* grad_outs, const int64_t* face_idx, const T* weights, const T* img,
                       const T* const* feats, const int* channels, int n_feat, int B, int F, int H, int W,
                       double multiplier, float eps, T* grad_img, T* const* grad_feats, void* stream,
                       bool zero_grad_img = true) {
  if (int e = check_dims(B, (int64_t)B * F, H, W)) return e;
  if (F < 0) return DIBR_ERR_INVALID_ARGUMENT;
  if (!weights && !((float)multiplier > 0.0f)) return DIBR_ERR_INVALID_ARGUMENT;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t nf = (size_t)B * F;
  FeatSet<const T> fin; FeatSet<T> gout;
  if (int e = make_featsets<T>(feats, grad_feats, channels, n_feat, nf > 0, fin, gout)) return e;
  FeatSet<const T> gin = fin;
  for (int t = 0; t < n_feat; ++t) {
    if (channels[t] > 0 && (int64_t)B * H * W > 0 && (!grad_outs || !grad_outs[t])) return DIBR_ERR_INVALID_ARGUMENT;
    gin.ptr[t] = grad_outs ? grad_outs[t] : nullptr;
  }
  if (nf > 0) {
    if (!grad_img) return DIBR_ERR_INVALID_ARGUMENT;
    if (zero_grad_img) CUDA_TRY(cudaMemsetAsync(grad_img, 0, nf * 6 * sizeof(T), s));
    for (int t = 0; t < n_feat; ++t)
      if (channels[t] > 0) CUDA_TRY(cudaMemsetAsync(grad_feats[t], 0, nf * 3 * channels[t] * sizeof(T), s));
  }
  if (B == 0 || H == 0 || W == 0 || nf == 0) return DIBR_OK;
  if (!face_idx || !img) return DIBR_ERR_INVALID_ARGUMENT;
  const Geometry geo = make_geometry(B, H, W);
  RasterBwdParams<T> bp;
  bp.grad_out = gin; bp.face_idx = face_idx; bp.weights = weights; bp.img = img; bp.feat = fin;
  bp.coord_mult = (T)multiplier; bp.pix_mult = (float)multiplier; bp.eps = eps; bp.F = F;
  bp.sx = (float)multiplier / (float)W; bp.sy = (float)multiplier / (float)H;      // = pix_scale(), IEEE fp32 division
  bp.grad_img = grad_img; bp.grad_feat = gout;
  {
    ProfScope ps(DIBR_KERNEL_RASTER_BWD, s);
    raster_bwd_kernel<T><<<dim3((unsigned)geo.tiles_x, (unsigned)geo.tiles_y, (unsigned)(B < 65535 ? B : 65535)), 256,
                           0, s>>>(bp, geo);
  }
  LAUNCH_CHECK("raster_bwd_kernel");
  return DIBR_OK;
}

// Threads per tile CTA of the fp32 soft-mask kernels: 128 (more independent tiles per SM hide the barrier-separated
// phases), except for DENSE scenes — many faces per tile, i.e. silhouette tiles with thousands of candidates — where
// 256 threads (2048 candidates per pass instead of 1024, twice the lanes per heavy tile) win by far: measured on B200,
// soft forward with 128 / 256 threads: c5 (61 faces per tile) 2.39 / 1.11 ms, c3 (24) 1.64 / 2.08 ms.
// DIBR_SOFT_NT (with DIBR_B200_TUNING=1) overrides the choice.
constexpr int SOFT_DENSE_FACES_PER_TILE = 48;
inline int soft_threads_default(int64_t faces_per_view, int tiles_per_view) {
  static const int forced = [] {
    const char* e = tuning_env("DIBR_SOFT_NT");
    const int v = e ? atoi(e) : 0;
    return (v == 64 || v == 128 || v == 256) ? v : 0;
  }();
  if (forced) return forced;
  return faces_per_view >= (int64_t)SOFT_DENSE_FACES_PER_TILE * tiles_per_view ? 256 : 128;
}

template <typename T, int MODE, int NT>
int launch_soft(const SoftParams<T>& p, const FaceLayout& fl, const Geometry& geo, const BinLists& L,
                int kernel_id, cudaStream_t s) {
  const size_t smem = SoftSmem<T, NT>::total;
  {
    // once per device and instantiation: not a stream operation, kept out of the per-call path so that the
    // entry points can be captured into CUDA graphs
    static std::atomic<unsigned long long> done{0};
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    const unsigned long long bit = 1ull << (dev & 63);
    if (!(done.load(std::memory_order_acquire) & bit)) {
      CUDA_TRY(cudaFuncSetAttribute(soft_mask_kernel<T, MODE, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)smem));
      done.fetch_or(bit, std::memory_order_release);
    }
  }
  ProfScope ps(kernel_id, s);
  soft_mask_kernel<T, MODE, NT><<<scheduled_grid(geo), NT, smem, s>>>(p, fl, geo, L);
  return DIBR_OK;
}

// persistent recompute of the tiles that overflowed the pair log (MODE_BWD)
template <typename T, int NT>
int launch_soft_overflow(const SoftParams<T>& p, const FaceLayout& fl, const Geometry& geo, const BinLists& L,
                         unsigned tiles, cudaStream_t s) {
  const size_t smem = SoftSmem<T, NT>::total;
  {
    static std::atomic<unsigned long long> done{0};
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    const unsigned long long bit = 1ull << (dev & 63);
    if (!(done.load(std::memory_order_acquire) & bit)) {
      CUDA_TRY(cudaFuncSetAttribute(soft_mask_overflow_kernel<T, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)smem));
      done.fetch_or(bit, std::memory_order_release);
    }
  }
  const unsigned grid = tiles < 296u ? tiles : 296u;          // 2 CTAs per SM
  soft_mask_overflow_kernel<T, NT><<<grid, NT, smem, s>>>(p, fl, geo, L);
  return DIBR_OK;
}

// Pair log of the soft-mask forward (dibr_soft_mask.cuh) and the FUSED workspace layout every soft-mask call uses:
//   [soft zero | raster zero (dibr_rasterization only) | log zero (always reserved) | soft rest | raster rest | log rest]
// The offsets do not depend on whether the log is in use (only the tail does), so a forward that had to go without a
// log is followed by a recompute backward on the same workspace.
struct LogLayout { size_t counts, flags, over, recs; unsigned cap_seg; bool ok; };
struct FusedLayout {
  WsLayout soft, rast;
  LogLayout log;
  bool has_rast;
  size_t unc;                      // per-tile coverage maps handed from the rasterizer to the soft mask (has_rast only)
  size_t zero_begin, zero_end, total_no_log, total;
};
FusedLayout fused_layout(int B, int64_t Ntot, bool with_raster, const Geometry& geo) {
  FusedLayout f;
  f.has_rast = with_raster;
  const int64_t P = (int64_t)geo.B * geo.H * geo.W;
  LogLayout& l = f.log;
  l.ok = P > 0 && P < ((int64_t)1 << 32) && Ntot < ((int64_t)1 << 29);   // pixel / face-view indices are logged as u32 / 29 bits
  int64_t cap = 2 * P + 65536;                            // measured: 0.5 - 1.3 pairs per pixel on configs 1-5
  static const long long forced = [] {                    // DIBR_SOFT_LOG_CAP=<entries>: tests of the overflow path
    const char* e = tuning_env("DIBR_SOFT_LOG_CAP");
    return e ? atoll(e) : 0ll;
  }();
  if (forced > 0) cap = forced;
  l.cap_seg = (unsigned)((cap + LOG_SEGMENTS - 1) / LOG_SEGMENTS);
  const size_t nt = (size_t)geo.B * geo.tiles_per_view;
  size_t o = 0;
  f.zero_begin = o;
  o = ws_zero_part(f.soft, B, geo, o);
  if (with_raster) o = ws_zero_part(f.rast, B, geo, o);
  l.counts = o; o = align_up(o + (LOG_SEGMENTS + 1) * sizeof(unsigned));     // + the overflow counter
  l.flags = o;  o = align_up(o + nt * sizeof(unsigned));
  f.zero_end = o;
  o = ws_rest_part(f.soft, B, Ntot, geo, o);
  if (with_raster) o = ws_rest_part(f.rast, B, Ntot, geo, o);
  f.unc = o;
  if (with_raster) o = align_up(o + nt * (TILE_PIX / 32) * sizeof(unsigned));
  f.total_no_log = o;
  l.over = o;   o = align_up(o + nt * sizeof(unsigned));
  l.recs = o;   o = align_up(o + (size_t)l.cap_seg * LOG_SEGMENTS * sizeof(uint4));   // 16-byte records, fp32 and fp64
  f.total = o;
  return f;
}

// The soft-mask kernels on prepared lists.  `use_log`: MODE_FWD appends its evaluated pairs to the log; MODE_BWD
// consumes it (log kernel + the normally idle overflow kernel) instead of re-deriving the k-lists.
template <typename T, int MODE>
int soft_launch_on_lists(SoftParams<T> p, const FaceLayout& fl, const Geometry& geo, const BinLists& L, void* ws,
                         const FusedLayout& fz, bool use_log, cudaStream_t s) {
  p.log_count = nullptr; p.over_count = nullptr; p.tile_flag = nullptr; p.over_list = nullptr; p.log = nullptr;
  p.log_cap = 0;
  {
    static const int sel = [] { const char* e = tuning_env("DIBR_SOFT_SEL"); return e ? atoi(e) : 0; }();
    p.sel_first = sel;                 // experiments: candidates taken by a tile's first pass (0 = default)
  }
  // backward from the log = two launches (logged pairs, then the normally idle recompute kernel): one scope
  ProfScope both(use_log && MODE == MODE_BWD ? DIBR_KERNEL_SOFT_BWD : -100, s);
  if (use_log) {
    unsigned char* base = static_cast<unsigned char*>(ws);
    p.log_count = reinterpret_cast<unsigned*>(base + fz.log.counts);
    p.over_count = p.log_count + LOG_SEGMENTS;
    p.tile_flag = reinterpret_cast<unsigned*>(base + fz.log.flags);
    p.over_list = reinterpret_cast<unsigned*>(base + fz.log.over);
    p.log = reinterpret_cast<uint4*>(base + fz.log.recs);
    p.log_cap = fz.log.cap_seg;
    if (MODE == MODE_BWD && fl.Ntot > 0) {     // the logged pairs first, then the (normally idle) recompute kernel
      soft_mask_bwd_log_kernel<T><<<dim3(LOG_CTAS_PER_SEGMENT, LOG_SEGMENTS), 256, 0, s>>>(p, fl, geo);
      g_launches.fetch_add(1, std::memory_order_relaxed);
    }
  }
  const unsigned grid = (unsigned)(geo.B * geo.tiles_per_view);
  const int kid = MODE == MODE_FWD ? DIBR_KERNEL_SOFT_FWD : DIBR_KERNEL_SOFT_BWD;
  // threads per CTA (one CTA per tile either way): fp32 uses small CTAs (more resident tiles per SM hide the
  // barrier-separated phases); DIBR_SOFT_NT=64|128|256 overrides for experiments.  fp64: 256.
  int nt = sizeof(T) == 4 ? soft_threads_default(geo.B > 0 ? fl.Ntot / geo.B : 0, geo.tiles_per_view) : 256;
  if (MODE == MODE_BWD && use_log) {
    if (sizeof(T) == 4) { if (int e = launch_soft_overflow<T, 128>(p, fl, geo, L, grid, s)) return e; }
    else { if (int e = launch_soft_overflow<T, 256>(p, fl, geo, L, grid, s)) return e; }
  } else if (sizeof(T) == 4 && nt == 64) {
    if (int e = launch_soft<T, MODE, 64>(p, fl, geo, L, kid, s)) return e;
  } else if (sizeof(T) == 4 && nt == 128) {
    if (int e = launch_soft<T, MODE, 128>(p, fl, geo, L, kid, s)) return e;
  } else {
    if (int e = launch_soft<T, MODE, 256>(p, fl, geo, L, kid, s)) return e;
  }
  LAUNCH_CHECK("soft_mask_kernel");
  return DIBR_OK;
}

// bins_ready: 0 = bin here; 1 = the forward call's bins are in `ws`; 2 = bins and pair log are; +4 = `ws` was laid out
// by the dibr_rasterization call (it also holds the rasterizer's lists).
template <typename T, int MODE>
int soft_mask_impl(const T* img, const T* bbox_in, const int64_t* face_idx, int B, int F, int H, int W,
                   float sigmainv, int knum, T coord_mult, int premultiplied, T margin, float pix_mult,
                   T* soft, T* prob, int64_t* close_idx, uint8_t* dist_type, const T* grad_soft, const T* soft_in,
                   T* grad_img, void* ws, size_t ws_bytes, void* stream, int bins_ready = 0, int keep_pairs = 0,
                   bool zero_grad = true) {
  if (int e = check_dims(B, (int64_t)B * F, H, W)) return e;
  if (F < 0 || knum < 0 || !(pix_mult > 0.0f)) return DIBR_ERR_INVALID_ARGUMENT;
  const bool side = prob || close_idx || dist_type;
  if (side && !(prob && close_idx && dist_type)) return DIBR_ERR_INVALID_ARGUMENT;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t nf = (size_t)B * F;
  if (MODE == MODE_BWD && nf > 0) {
    if (!grad_img) return DIBR_ERR_INVALID_ARGUMENT;
    if (zero_grad) CUDA_TRY(cudaMemsetAsync(grad_img, 0, nf * 6 * sizeof(T), s));
  }
  if (B == 0 || H == 0 || W == 0) return DIBR_OK;
  if (!face_idx || (nf > 0 && !img)) return DIBR_ERR_INVALID_ARGUMENT;
  if (MODE == MODE_FWD && !soft) return DIBR_ERR_INVALID_ARGUMENT;
  if (MODE == MODE_BWD && (!grad_soft || !soft_in)) return DIBR_ERR_INVALID_ARGUMENT;
  const Geometry geo = make_geometry(B, H, W);
  FaceLayout fl; fl.first_idx = nullptr; fl.F = F; fl.Ntot = (int64_t)nf; fl.B = B;
  const FusedLayout fz = fused_layout(B, fl.Ntot, (bins_ready & 4) != 0, geo);
  if (!ws || ws_bytes < fz.total_no_log) return DIBR_ERR_WORKSPACE;
  // The pair log is used when it was asked for (forward: keep_pairs; backward: bins_ready & 2), the pixel indices
  // fit 32 bits and the workspace holds it — the same test on both sides.
  const bool use_log = ((MODE == MODE_FWD) ? (keep_pairs != 0 && !side) : ((bins_ready & 2) != 0)) && fz.log.ok &&
                       ws_bytes >= fz.total;
  BinLists L = carve(ws, fz.soft);
  if (!(bins_ready & 3)) {                  // otherwise: the forward call's bins, still in `ws` (only read here)
    CUDA_TRY(cudaMemsetAsync(static_cast<unsigned char*>(ws) + fz.zero_begin, 0, fz.zero_end - fz.zero_begin, s));
    SetupParams<T> sp;
    sp.img = img; sp.bbox_in = bbox_in; sp.valid = nullptr; sp.normals_z = nullptr; sp.coord_mult = coord_mult;
    sp.premultiplied = premultiplied; sp.margin = margin; sp.pix_mult = pix_mult;
    BinLists2 LL; LL.l[0] = L; LL.l[1] = L;
    const unsigned nb = binning_blocks(fl, geo, false);
    if (nb > 0) {
      { ProfScope ps(DIBR_KERNEL_SETUP_COUNT, s); setup_count_kernel<T, false, false><<<nb, 256, 0, s>>>(sp, (T)0, fl, geo, LL, PrepareParams<T>(), MeshOut<T>()); }
      LAUNCH_CHECK("setup_count_kernel");
    }
    { ProfScope ps(DIBR_KERNEL_SCAN, s); scan_kernel<<<scan_grid(geo, 1), SCAN_CHUNK, 0, s>>>(geo, LL); }
    LAUNCH_CHECK("scan_kernel");
    { ProfScope ps(DIBR_KERNEL_FILL, s); fill_kernel<false><<<binning_blocks(fl, geo, true), 256, 0, s>>>(fl, geo, LL); }
    LAUNCH_CHECK("fill_kernel");
  }
  SoftParams<T> p;
  p.img = img; p.coord_mult = coord_mult; p.premultiplied = premultiplied; p.pix_mult = pix_mult;
  p.sigmainv = sigmainv; p.F = F; p.knum = knum; p.face_idx = face_idx; p.soft = soft; p.prob = prob;
  p.close_idx = close_idx; p.dist_type = dist_type; p.grad_soft = grad_soft; p.soft_in = soft_in;
  p.grad_img = grad_img; p.tile_unc = nullptr;
  return soft_launch_on_lists<T, MODE>(p, fl, geo, L, ws, fz, use_log, s);
}

// dibr_rasterization (dibr.py:119-209) = rasterize(valid = face_normals_z >= 0) + dibr_soft_mask in ONE call: both
// list sets from one read of the coordinates (setup / scan / fill once), then the two per-tile kernels.
template <typename T>
int dibr_rasterization_fwd_impl(const T* z, const T* img, const T* const* feats, const int* channels, int n_feat,
                                const T* normals_z, const uint8_t* valid, int B, int F, int H, int W,
                                double multiplier, float eps, float sigmainv, double boxlen, int knum,
                                T* const* out_feats, int64_t* out_idx, T* soft, void* ws, size_t ws_bytes,
                                int keep_pairs, void* stream, const PrepareParams<T>* mesh = nullptr) {
  if (F < 0 || knum < 0) return DIBR_ERR_INVALID_ARGUMENT;
  const int64_t Ntot = (int64_t)B * F;
  if (int e = check_dims(B, Ntot, H, W)) return e;
  const float pix_mult = (float)multiplier;
  if (!(pix_mult > 0.0f)) return DIBR_ERR_INVALID_ARGUMENT;
  if (B == 0 || H == 0 || W == 0) return DIBR_OK;
  if ((Ntot > 0 && (!z || !img)) || !out_idx || !soft) return DIBR_ERR_INVALID_ARGUMENT;
  FeatSet<const T> fin; FeatSet<T> fout;
  if (int e = make_featsets<T>(feats, out_feats, channels, n_feat, Ntot > 0, fin, fout)) return e;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const Geometry geo = make_geometry(B, H, W);
  FaceLayout fl; fl.first_idx = nullptr; fl.F = F; fl.Ntot = Ntot; fl.B = B;
  const FusedLayout fz = fused_layout(B, Ntot, true, geo);
  if (!ws || ws_bytes < fz.total_no_log) return DIBR_ERR_WORKSPACE;
  const bool use_log = keep_pairs != 0 && fz.log.ok && ws_bytes >= fz.total;
  BinLists2 LL; LL.l[0] = carve(ws, fz.rast); LL.l[1] = carve(ws, fz.soft);
  CUDA_TRY(cudaMemsetAsync(static_cast<unsigned char*>(ws) + fz.zero_begin, 0, fz.zero_end - fz.zero_begin, s));
  SetupParams<T> sp;
  sp.img = img; sp.bbox_in = nullptr; sp.valid = valid; sp.normals_z = normals_z; sp.coord_mult = (T)multiplier;
  sp.premultiplied = 0; sp.margin = (T)0; sp.pix_mult = pix_mult;
  const unsigned nb = binning_blocks(fl, geo, false);
  if (nb > 0) {
    { ProfScope ps(DIBR_KERNEL_SETUP_COUNT, s);
      if (mesh) {         // prepare_vertices folded into K0: `img` / `z` are OUTPUTS of this kernel (const cast: the
        MeshOut<T> mo;    // caller's buffers, handed over as outputs by dibr_rasterization_from_mesh_fwd)
        mo.img = const_cast<T*>(img); mo.z = const_cast<T*>(z);
        setup_count_kernel<T, true, true><<<nb, 256, 0, s>>>(sp, (T)(boxlen * multiplier), fl, geo, LL, *mesh, mo);
      } else {
        setup_count_kernel<T, true, false><<<nb, 256, 0, s>>>(sp, (T)(boxlen * multiplier), fl, geo, LL,
                                                              PrepareParams<T>(), MeshOut<T>());
      } }
    LAUNCH_CHECK("setup_count_kernel");
  }
  { ProfScope ps(DIBR_KERNEL_SCAN, s); scan_kernel<<<scan_grid(geo, 2), SCAN_CHUNK, 0, s>>>(geo, LL); }
  LAUNCH_CHECK("scan_kernel");
  { ProfScope ps(DIBR_KERNEL_FILL, s); fill_kernel<true><<<binning_blocks(fl, geo, true), 256, 0, s>>>(fl, geo, LL); }
  LAUNCH_CHECK("fill_kernel");
  RasterFwdParams<T> rp;
  rp.z = z; rp.img = img; rp.feat = fin; rp.coord_mult = (T)multiplier; rp.premultiplied = 0;
  rp.pix_mult = pix_mult; rp.eps = eps; rp.out_feat = fout; rp.out_idx = out_idx; rp.out_w = nullptr;
  // fp32: the rasterizer's epilogue hands every tile's coverage map (and the provisional soft values) to the
  // soft-mask kernel; the fp64 gather kernel does not, and the soft-mask kernel then reads face_idx as usual
  unsigned* const tile_unc = sizeof(T) == 4 ? reinterpret_cast<unsigned*>(static_cast<unsigned char*>(ws) + fz.unc)
                                            : nullptr;
  rp.soft_out = tile_unc ? soft : nullptr; rp.tile_unc = tile_unc;
  {
    ProfScope ps(DIBR_KERNEL_RASTER_FWD, s);
    launch_raster_fwd(rp, fl, geo, LL.l[0], s);
  }
  LAUNCH_CHECK("raster_fwd_kernel");
  SoftParams<T> p;
  p.img = img; p.coord_mult = (T)multiplier; p.premultiplied = 0; p.pix_mult = pix_mult;
  p.sigmainv = sigmainv; p.F = F; p.knum = knum; p.face_idx = out_idx; p.soft = soft; p.prob = nullptr;
  p.close_idx = nullptr; p.dist_type = nullptr; p.grad_soft = nullptr; p.soft_in = nullptr; p.grad_img = nullptr;
  p.tile_unc = tile_unc;
  return soft_launch_on_lists<T, MODE_FWD>(p, fl, geo, LL.l[1], ws, fz, use_log, s);
}

// Backward of both outputs of dibr_rasterization into ONE grad_img buffer (zero-filled once; the two kernels'
// red.global.add accumulate into it) + the feature gradients.  Either upstream gradient may be absent.
template <typename T>
int dibr_rasterization_bwd_impl(const T* const* grad_outs, const T* grad_soft, const T* soft,
                                const int64_t* face_idx, const T* img, const T* const* feats, const int* channels,
                                int n_feat, int B, int F, int H, int W, double multiplier, float eps, float sigmainv,
                                double boxlen, int knum, T* grad_img, T* const* grad_feats, void* ws, size_t ws_bytes,
                                int bins_ready, void* stream) {
  if (F < 0 || knum < 0) return DIBR_ERR_INVALID_ARGUMENT;
  if (int e = check_dims(B, (int64_t)B * F, H, W)) return e;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t nf = (size_t)B * F;
  if (nf > 0) {
    if (!grad_img) return DIBR_ERR_INVALID_ARGUMENT;
    CUDA_TRY(cudaMemsetAsync(grad_img, 0, nf * 6 * sizeof(T), s));
  }
  bool any_out = false;
  for (int t = 0; t < n_feat; ++t) any_out = any_out || (grad_outs && grad_outs[t]);
  if (any_out || n_feat > 0) {
    // feature gradients are zero-filled by rasterize_bwd_impl; with no upstream feature gradient they stay zero
    if (any_out) {
      if (int e = rasterize_bwd_impl<T>(grad_outs, face_idx, nullptr, img, feats, channels, n_feat, B, F, H, W,
                                        multiplier, eps, grad_img, grad_feats, stream, false))
        return e;
    } else if (nf > 0 && grad_feats) {
      for (int t = 0; t < n_feat; ++t)
        if (channels[t] > 0 && grad_feats[t]) CUDA_TRY(cudaMemsetAsync(grad_feats[t], 0, nf * 3 * channels[t] * sizeof(T), s));
    }
  }
  if (grad_soft) {
    if (int e = soft_mask_impl<T, MODE_BWD>(img, nullptr, face_idx, B, F, H, W, sigmainv, knum, (T)multiplier, 0,
                                            (T)(boxlen * multiplier), (float)multiplier, nullptr, nullptr, nullptr,
                                            nullptr, grad_soft, soft, grad_img, ws, ws_bytes, stream, bins_ready, 0,
                                            false))
      return e;
  }
  return DIBR_OK;
}

template <typename T>
int soft_mask_bwd_sidebuf_impl(const T* grad_soft, const T* soft, const int64_t* face_idx, const T* prob,
                               const int64_t* close_idx, const uint8_t* dist_type, const T* img_mult, int B, int F,
                               int H, int W, int knum, float sigmainv, float multiplier, T* grad_img, void* stream) {
  if (int e = check_dims(B, (int64_t)B * F, H, W)) return e;
  if (F < 0 || knum < 0) return DIBR_ERR_INVALID_ARGUMENT;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t nf = (size_t)B * F;
  if (nf > 0) {
    if (!grad_img) return DIBR_ERR_INVALID_ARGUMENT;
    CUDA_TRY(cudaMemsetAsync(grad_img, 0, nf * 6 * sizeof(T), s));
  }
  const int64_t P = (int64_t)B * H * W;
  if (P == 0 || nf == 0 || knum == 0) return DIBR_OK;
  if (!grad_soft || !soft || !face_idx || !prob || !close_idx || !dist_type || !img_mult)
    return DIBR_ERR_INVALID_ARGUMENT;
  soft_mask_bwd_sidebuf_kernel<T><<<(unsigned)((P + 255) / 256), 256, 0, s>>>(
      grad_soft, soft, face_idx, prob, close_idx, dist_type, img_mult, grad_img, B, H, W, F, knum, sigmainv,
      multiplier);
  LAUNCH_CHECK("soft_mask_bwd_sidebuf_kernel");
  return DIBR_OK;
}

// prepare_vertices (utils.py:128-175): one kernel each way; grad_vertices is zero-filled here.
template <typename T>
int prepare_impl(bool backward, const T* g_cam, const T* g_img, const T* g_nrm, const T* vertices,
                 const int64_t* faces, const T* proj, const T* rot, const T* trans, const T* transform, int B,
                 int Bv, int V, int F, T* face_cam, T* face_img, T* face_nrm, T* g_vertices, T* g_proj, T* g_rot,
                 T* g_trans, T* g_transform, void* stream) {
  if (B < 0 || V < 0 || F < 0 || (Bv != 1 && Bv != B)) return DIBR_ERR_INVALID_ARGUMENT;
  if ((int64_t)B * F >= (int64_t)1 << 40 || (int64_t)Bv * V >= (int64_t)1 << 40) return DIBR_ERR_TOO_LARGE;
  const bool by_matrix = transform != nullptr;
  if (by_matrix == (rot != nullptr || trans != nullptr)) return DIBR_ERR_INVALID_ARGUMENT;   // exactly one camera form
  if (!by_matrix && (!rot || !trans)) return DIBR_ERR_INVALID_ARGUMENT;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t n = (int64_t)B * F;
  const bool cam_grad = g_proj || g_rot || g_trans || g_transform;
  if (backward) {
    if ((by_matrix && (g_rot || g_trans)) || (!by_matrix && g_transform)) return DIBR_ERR_INVALID_ARGUMENT;
    if (!g_vertices && !cam_grad) return DIBR_ERR_INVALID_ARGUMENT;
    if (g_vertices && (int64_t)Bv * V > 0) CUDA_TRY(cudaMemsetAsync(g_vertices, 0, (size_t)Bv * V * 3 * sizeof(T), s));
    if (g_proj) CUDA_TRY(cudaMemsetAsync(g_proj, 0, 3 * sizeof(T), s));
    if (g_rot && B > 0) CUDA_TRY(cudaMemsetAsync(g_rot, 0, (size_t)B * 9 * sizeof(T), s));
    if (g_trans && B > 0) CUDA_TRY(cudaMemsetAsync(g_trans, 0, (size_t)B * 3 * sizeof(T), s));
    if (g_transform && B > 0) CUDA_TRY(cudaMemsetAsync(g_transform, 0, (size_t)B * 12 * sizeof(T), s));
  } else if (n > 0 && (!face_cam || !face_img || !face_nrm)) {
    return DIBR_ERR_INVALID_ARGUMENT;
  }
  if (n == 0) return DIBR_OK;
  if (!vertices || !faces || !proj) return DIBR_ERR_INVALID_ARGUMENT;
  PrepareParams<T> pp;
  pp.vertices = vertices; pp.faces = faces; pp.proj = proj; pp.rot = rot; pp.trans = trans;
  pp.transform = transform; pp.B = B; pp.Bv = Bv; pp.V = V; pp.F = F;
  pp.face_cam = face_cam; pp.face_img = face_img; pp.face_nrm = face_nrm;
  pp.g_cam = g_cam; pp.g_img = g_img; pp.g_nrm = g_nrm; pp.g_vertices = g_vertices;
  pp.g_proj = g_proj; pp.g_rot = g_rot; pp.g_trans = g_trans; pp.g_transform = g_transform;
  const unsigned nb = (unsigned)((n + 255) / 256);
  if (backward) {
    if (!g_cam && !g_img && !g_nrm) return DIBR_OK;                 // all-zero upstream: the memset is the answer
    { ProfScope ps(DIBR_KERNEL_PREPARE_BWD, s);
      if (cam_grad) {
        prepare_vertices_bwd_kernel<T, true><<<nb, 256, 0, s>>>(pp);
      } else if (Bv == 1 && B >= 4 && g_vertices && F > 0) {
        // one mesh, many cameras: the face's gradient summed over a chunk of views before it is scattered
        const int vpt = B <= 32 ? 4 : 8;
        prepare_vertices_bwd_shared_kernel<T><<<dim3((unsigned)((F + 255) / 256), (unsigned)((B + vpt - 1) / vpt)), 256, 0,
                                                s>>>(pp, vpt);
      } else {
        prepare_vertices_bwd_kernel<T, false><<<nb, 256, 0, s>>>(pp);
      } }
    LAUNCH_CHECK("prepare_vertices_bwd_kernel");
  } else {
    { ProfScope ps(DIBR_KERNEL_PREPARE_FWD, s); prepare_vertices_fwd_kernel<T><<<nb, 256, 0, s>>>(pp); }
    LAUNCH_CHECK("prepare_vertices_fwd_kernel");
  }
  return DIBR_OK;
}

// Validation shared by the mesh entry points (same rules as prepare_impl).
template <typename T>
int mesh_params(const T* vertices, const int64_t* faces, const T* proj, const T* rot, const T* trans,
                const T* transform, int B, int Bv, int V, int F, PrepareParams<T>& pp) {
  if (B < 0 || V < 0 || F < 0 || (Bv != 1 && Bv != B)) return DIBR_ERR_INVALID_ARGUMENT;
  if ((int64_t)Bv * V >= (int64_t)1 << 40) return DIBR_ERR_TOO_LARGE;
  const bool by_matrix = transform != nullptr;
  if (by_matrix == (rot != nullptr || trans != nullptr)) return DIBR_ERR_INVALID_ARGUMENT;   // exactly one camera form
  if (!by_matrix && (!rot || !trans)) return DIBR_ERR_INVALID_ARGUMENT;
  if ((int64_t)B * F > 0 && (!vertices || !faces || !proj)) return DIBR_ERR_INVALID_ARGUMENT;
  pp = PrepareParams<T>();
  pp.vertices = vertices; pp.faces = faces; pp.proj = proj; pp.rot = rot; pp.trans = trans;
  pp.transform = transform; pp.B = B; pp.Bv = Bv; pp.V = V; pp.F = F;
  return DIBR_OK;
}

// dibr_rasterization straight from the mesh: prepare_vertices (utils.py:128-175) + dibr_rasterization (dibr.py:119-209)
// as 5 launches — K0 projects the faces itself (setup_count_kernel<T, true, true>) and writes face_vertices_image /
// face_vertices_z for the per-tile kernels and the backward.
template <typename T>
int from_mesh_fwd_impl(const T* vertices, const int64_t* faces, const T* proj, const T* rot, const T* trans,
                       const T* transform, int Bv, int V, const T* const* feats, const int* channels, int n_feat,
                       int B, int F, int H, int W, double multiplier, float eps, float sigmainv, double boxlen,
                       int knum, T* face_img, T* face_z, T* const* out_feats, int64_t* out_idx, T* soft, void* ws,
                       size_t ws_bytes, int keep_pairs, void* stream) {
  PrepareParams<T> pp;
  if (int e = mesh_params<T>(vertices, faces, proj, rot, trans, transform, B, Bv, V, F, pp)) return e;
  if ((int64_t)B * F > 0 && (!face_img || !face_z)) return DIBR_ERR_INVALID_ARGUMENT;
  return dibr_rasterization_fwd_impl<T>(face_z, face_img, feats, channels, n_feat, nullptr, nullptr, B, F, H, W,
                                        multiplier, eps, sigmainv, boxlen, knum, out_feats, out_idx, soft, ws, ws_bytes,
                                        keep_pairs, stream, &pp);
}

// Backward of the mesh entry: image-plane gradients of both outputs into the caller's scratch (B,F,3,2), then
// prepare_vertices' backward folds them into grad_vertices (+ optional camera gradients) — one call, the (B,F,3,2)
// gradient never leaves the library's hands.
template <typename T>
int from_mesh_bwd_impl(const T* const* grad_outs, const T* grad_soft, const T* soft, const int64_t* face_idx,
                       const T* face_img, const T* const* feats, const int* channels, int n_feat, const T* vertices,
                       const int64_t* faces, const T* proj, const T* rot, const T* trans, const T* transform, int Bv,
                       int V, int B, int F, int H, int W, double multiplier, float eps, float sigmainv, double boxlen,
                       int knum, T* grad_img_scratch, T* const* grad_feats, T* g_vertices, T* g_proj, T* g_rot,
                       T* g_trans, T* g_transform, void* ws, size_t ws_bytes, int bins_ready, void* stream) {
  if (int e = dibr_rasterization_bwd_impl<T>(grad_outs, grad_soft, soft, face_idx, face_img, feats, channels, n_feat, B,
                                             F, H, W, multiplier, eps, sigmainv, boxlen, knum, grad_img_scratch,
                                             grad_feats, ws, ws_bytes, bins_ready, stream))
    return e;
  return prepare_impl<T>(true, nullptr, grad_img_scratch, nullptr, vertices, faces, proj, rot, trans, transform, B, Bv,
                         V, F, nullptr, nullptr, nullptr, g_vertices, g_proj, g_rot, g_trans, g_transform, stream);
}

// texture_mapping (utils.py:23-76) [+ clamp(image * mask, 0, 1)]: one kernel each way.
template <typename T>
int texture_impl(bool backward, const T* g_out, const T* uv, const T* tex, const T* mask, int B, int64_t N, int Bt,
                 int C, int Ht, int Wt, int mode, T* out, T* g_uv, T* g_tex, T* g_mask, void* stream) {
  if (B < 0 || N < 0 || C < 0 || Ht < 0 || Wt < 0 || (Bt != 1 && Bt != B)) return DIBR_ERR_INVALID_ARGUMENT;
  if (mode != TEX_NEAREST && mode != TEX_BILINEAR) return DIBR_ERR_INVALID_ARGUMENT;
  if ((int64_t)B * N >= (int64_t)1 << 40 || Ht > 32768 || Wt > 32768) return DIBR_ERR_TOO_LARGE;
  if (g_mask && !mask) return DIBR_ERR_INVALID_ARGUMENT;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t n = (int64_t)B * N;
  const size_t tex_elems = (size_t)Bt * C * Ht * Wt;
  if (backward && g_tex && tex_elems > 0) CUDA_TRY(cudaMemsetAsync(g_tex, 0, tex_elems * sizeof(T), s));
  if (n == 0 || C == 0) return DIBR_OK;
  if (Ht == 0 || Wt == 0) return DIBR_ERR_INVALID_ARGUMENT;       // grid_sample rejects empty spatial sizes
  if (!uv || !tex || (backward ? !g_out : !out)) return DIBR_ERR_INVALID_ARGUMENT;
  TextureParams<T> tp;
  tp.uv = uv; tp.tex = tex; tp.mask = mask; tp.B = B; tp.Bt = Bt; tp.C = C; tp.Ht = Ht; tp.Wt = Wt; tp.N = N;
  tp.mode = mode; tp.out = out; tp.g_out = g_out; tp.g_uv = g_uv; tp.g_tex = g_tex; tp.g_mask = g_mask;
  const unsigned nb = (unsigned)((n + 255) / 256);
  if (backward) {
    { ProfScope ps(DIBR_KERNEL_TEXTURE_BWD, s); texture_mapping_bwd_kernel<T><<<nb, 256, 0, s>>>(tp); }
    LAUNCH_CHECK("texture_mapping_bwd_kernel");
  } else {
    { ProfScope ps(DIBR_KERNEL_TEXTURE_FWD, s); texture_mapping_fwd_kernel<T><<<nb, 256, 0, s>>>(tp); }
    LAUNCH_CHECK("texture_mapping_fwd_kernel");
  }
  return DIBR_OK;
}

// mask_iou (kaolin/metrics/render.py:18-40)
inline int iou_chunks(int64_t P) { return (int)((P + IOU_CHUNK - 1) / IOU_CHUNK); }
size_t iou_ws_bytes(int B, int64_t P, int elem) {
  if (B <= 0 || P <= 0) return 256;
  return align_up((size_t)B * iou_chunks(P) * 2 * (size_t)elem);
}
template <typename T>
int iou_fwd_impl(const T* lhs, const T* rhs, int B, int64_t P, T* loss, T* sums, void* ws, size_t ws_bytes,
                 void* stream) {
  if (B < 0 || P < 0 || !loss) return DIBR_ERR_INVALID_ARGUMENT;
  if (B > 65535 || P >= (int64_t)1 << 40) return DIBR_ERR_TOO_LARGE;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int chunks = iou_chunks(P);
  if (B > 0 && (!sums || (P > 0 && (!lhs || !rhs)))) return DIBR_ERR_INVALID_ARGUMENT;
  if (B > 0 && P > 0) {
    if (!ws || ws_bytes < iou_ws_bytes(B, P, (int)sizeof(T))) return DIBR_ERR_WORKSPACE;
    { ProfScope ps(DIBR_KERNEL_IOU_FWD, s);
      mask_iou_partial_kernel<T><<<dim3((unsigned)chunks, (unsigned)B), 256, 0, s>>>(lhs, rhs, P, chunks,
                                                                                      static_cast<T*>(ws)); }
    LAUNCH_CHECK("mask_iou_partial_kernel");
  }
  // B == 0: mean of an empty tensor is NaN in torch -> 1 - NaN; the finish kernel divides 0 by 0 likewise
  mask_iou_finish_kernel<T><<<1, 256, 0, s>>>(static_cast<const T*>(ws), B, P > 0 ? chunks : 0, sums, loss);
  LAUNCH_CHECK("mask_iou_finish_kernel");
  return DIBR_OK;
}
template <typename T>
int iou_bwd_impl(const T* g_loss, const T* lhs, const T* rhs, const T* sums, int B, int64_t P, T* g_lhs, T* g_rhs,
                 void* stream) {
  if (B < 0 || P < 0) return DIBR_ERR_INVALID_ARGUMENT;
  if (B > 65535 || P >= (int64_t)1 << 40) return DIBR_ERR_TOO_LARGE;
  if (B == 0 || P == 0 || (!g_lhs && !g_rhs)) return DIBR_OK;
  if (!g_loss || !lhs || !rhs || !sums) return DIBR_ERR_INVALID_ARGUMENT;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  { ProfScope ps(DIBR_KERNEL_IOU_BWD, s);
    mask_iou_bwd_kernel<T><<<dim3((unsigned)iou_chunks(P), (unsigned)B), 256, 0, s>>>(lhs, rhs, sums, g_loss, P,
                                                                                       B, g_lhs, g_rhs); }
  LAUNCH_CHECK("mask_iou_bwd_kernel");
  return DIBR_OK;
}

// image L1 loss (examples/tutorial/ian_dibr.py:260): mean |a - b| over n elements
inline int l1_chunks(int64_t n) { return (int)((n + L1_CHUNK - 1) / L1_CHUNK); }
template <typename T>
int l1_fwd_impl(const T* a, const T* b, int64_t n, T* loss, void* ws, size_t ws_bytes, void* stream) {
  if (n < 0 || !loss) return DIBR_ERR_INVALID_ARGUMENT;
  if (n >= (int64_t)1 << 44) return DIBR_ERR_TOO_LARGE;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int chunks = l1_chunks(n);
  if (n > 0) {
    if (!a || !b) return DIBR_ERR_INVALID_ARGUMENT;
    if (!ws || ws_bytes < align_up((size_t)chunks * sizeof(T))) return DIBR_ERR_WORKSPACE;
    { ProfScope ps(DIBR_KERNEL_L1_FWD, s);
      image_l1_partial_kernel<T><<<(unsigned)chunks, 256, 0, s>>>(a, b, n, static_cast<T*>(ws)); }
    LAUNCH_CHECK("image_l1_partial_kernel");
  }
  image_l1_finish_kernel<T><<<1, 256, 0, s>>>(static_cast<const T*>(ws), chunks, n, loss);
  LAUNCH_CHECK("image_l1_finish_kernel");
  return DIBR_OK;
}
template <typename T>
int l1_bwd_impl(const T* g_loss, const T* a, const T* b, int64_t n, T* g_a, T* g_b, void* stream) {
  if (n < 0) return DIBR_ERR_INVALID_ARGUMENT;
  if (n >= (int64_t)1 << 44) return DIBR_ERR_TOO_LARGE;
  if (n == 0 || (!g_a && !g_b)) return DIBR_OK;
  if (!g_loss || !a || !b) return DIBR_ERR_INVALID_ARGUMENT;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  { ProfScope ps(DIBR_KERNEL_L1_BWD, s);
    image_l1_bwd_kernel<T><<<(unsigned)l1_chunks(n), 256, 0, s>>>(a, b, g_loss, n, g_a, g_b); }
  LAUNCH_CHECK("image_l1_bwd_kernel");
  return DIBR_OK;
}

// Adam step for one parameter tensor (the fitting loop's optimiser; see dibr_losses.cuh).  state: step (1 float) and
// arrivals (1 unsigned, zero-initialised by the caller once) live on the device.
template <typename T>
int adam_impl(T* param, const T* grad, T* exp_avg, T* exp_avg_sq, float* step, unsigned* arrivals, int64_t n, double lr,
              double beta1, double beta2, double eps, void* stream) {
  if (n < 0) return DIBR_ERR_INVALID_ARGUMENT;
  if (n >= (int64_t)1 << 38) return DIBR_ERR_TOO_LARGE;
  if (!step || !arrivals || (n > 0 && (!param || !grad || !exp_avg || !exp_avg_sq))) return DIBR_ERR_INVALID_ARGUMENT;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const unsigned nb = (unsigned)((n + 255) / 256);
  adam_step_kernel<T><<<nb > 0 ? nb : 1, 256, 0, s>>>(param, grad, exp_avg, exp_avg_sq, step, arrivals, n, lr, beta1,
                                                      beta2, eps);
  LAUNCH_CHECK("adam_step_kernel");
  return DIBR_OK;
}

// ---- allocation-free variant of mask_iou forward (no counterpart in the reference; added
// for callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc).  Its scratch — the per-chunk partial sums of the two-stage reduction — lives in this file-scope
// device array instead of a caller-provided workspace.  Every GPU that loads the library holds these
// DIBR_NOALLOC_SCRATCH_BYTES (1 MiB) until the process exits.
//   * In HOST code `g_noalloc_scratch` is only a host-side stand-in for the symbol, NOT a device address: the address
//     is taken once per device, with that device current, by cudaGetSymbolAddress below, and that pointer is what the
//     kernels get.
//   * All calls on a device share the array and it keeps the last call's contents.  Nothing in it has to start at
//     zero: mask_iou_partial_kernel writes every (view, chunk) slot that mask_iou_finish_kernel reads, in the same
//     call.  Calls that overlap on the GPU (different streams, no ordering between them) RACE on it.
constexpr size_t NOALLOC_SCRATCH_BYTES = DIBR_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];
constexpr int NOALLOC_MAX_DEVICES = 64;
std::atomic<void*> g_noalloc_ptr[NOALLOC_MAX_DEVICES];

// device address of the scratch array on the CURRENT device (set-up: once per device; afterwards one atomic load)
int noalloc_scratch_ptr(void** out) {
  int dev = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return DIBR_ERR_INVALID_ARGUMENT;
  void* p = g_noalloc_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));      // `dev` is current: its copy of the array
    if (!p) return cuda_fail(cudaErrorInvalidSymbol, "cudaGetSymbolAddress(g_noalloc_scratch)");
    g_noalloc_ptr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return DIBR_OK;
}

template <typename T>
int iou_fwd_noalloc_impl(const T* lhs, const T* rhs, int B, int64_t P, T* loss, T* sums, void* stream) {
  if (B < 0 || P < 0 || !loss) return DIBR_ERR_INVALID_ARGUMENT;
  if (B > 65535 || P >= (int64_t)1 << 40) return DIBR_ERR_TOO_LARGE;
  const size_t need = iou_ws_bytes(B, P, (int)sizeof(T));
  if (need > NOALLOC_SCRATCH_BYTES) return DIBR_ERR_SCRATCH_LIMIT;     // before any CUDA call
  void* scratch = nullptr;
  if (int e = noalloc_scratch_ptr(&scratch)) return e;
  // the existing forward on the built-in scratch: two kernel launches on `stream`, no allocation of any kind
  return iou_fwd_impl<T>(lhs, rhs, B, P, loss, sums, scratch, NOALLOC_SCRATCH_BYTES, stream);
}

// deftet_sparse_render (deftet.py:269-417).  Grid resolution: about 8 faces per cell, G in [1, 512].
inline int deftet_grid(int F) {
  int g = (int)std::ceil(std::sqrt((double)(F > 0 ? F : 1) / 8.0));
  return g < 1 ? 1 : (g > 512 ? 512 : g);
}
inline int deftet_cap(int knum) {
  int c = 64;
  while (c < knum + 32) c <<= 1;
  return c;
}
constexpr int DEFTET_MAX_KNUM = 4064;

template <typename T>
int deftet_fwd_impl(const T* z, const T* img, const T* bbox, const T* pix, const T* ranges, const T* const* feats,
                    const int* channels, int n_feat, int B, int F, int P, int knum, float eps, T* const* outs,
                    int64_t* face_idx, T* weights, int64_t* idx_unsorted, T* depth, T* w0, T* w1, void* ws,
                    size_t ws_bytes, void* stream) {
  if (B < 0 || F < 0 || P < 0 || knum < 0) return DIBR_ERR_INVALID_ARGUMENT;
  if (knum > DEFTET_MAX_KNUM) return DIBR_ERR_TOO_LARGE;
  const int64_t Ntot = (int64_t)B * F;
  if (Ntot >= (int64_t)1 << 31 || (int64_t)B * P * (knum > 0 ? knum : 1) >= (int64_t)1 << 40) return DIBR_ERR_TOO_LARGE;
  if ((int64_t)B * P == 0 || knum == 0) return DIBR_OK;
  const bool native = idx_unsorted != nullptr;
  if (native ? (!depth || !w0 || !w1) : !face_idx) return DIBR_ERR_INVALID_ARGUMENT;
  if (!pix || !ranges || (Ntot > 0 && (!z || !img))) return DIBR_ERR_INVALID_ARGUMENT;
  FeatSet<const T> fin; FeatSet<T> fout;
  fin.n = fout.n = 0;
  if (!native) { if (int e = make_featsets<T>(feats, outs, channels, n_feat, Ntot > 0, fin, fout)) return e; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int G = deftet_grid(F);
  const Geometry geo = make_geometry(B, G * TILE, G * TILE);
  FaceLayout fl; fl.first_idx = nullptr; fl.F = F; fl.Ntot = Ntot; fl.B = B;
  const WsLayout w = ws_layout(B, Ntot, geo);
  if (!ws || ws_bytes < w.total) return DIBR_ERR_WORKSPACE;
  BinLists L = carve(ws, w);
  CUDA_TRY(cudaMemsetAsync(static_cast<unsigned char*>(ws) + w.zero_begin, 0, w.zero_end - w.zero_begin, s));
  const unsigned nb = (unsigned)((Ntot + 255) / 256);
  if (nb > 0) {
    deftet_setup_count_kernel<T><<<nb, 256, 0, s>>>(img, bbox, fl, geo, L, G);
    LAUNCH_CHECK("deftet_setup_count_kernel");
  }
  BinLists2 LL; LL.l[0] = L; LL.l[1] = L;
  scan_kernel<<<scan_grid(geo, 1), SCAN_CHUNK, 0, s>>>(geo, LL);
  LAUNCH_CHECK("scan_kernel");
  if (nb > 0) {
    fill_kernel<false><<<binning_blocks(fl, geo, true), 256, 0, s>>>(fl, geo, LL);
    LAUNCH_CHECK("fill_kernel");
  }
  DeftetParams<T> dp;
  dp.z = z; dp.img = img; dp.bbox = bbox; dp.pix = pix; dp.ranges = ranges;
  dp.B = B; dp.F = F; dp.P = P; dp.knum = knum; dp.grid = G; dp.eps = eps; dp.cap = deftet_cap(knum);
  dp.face_idx_unsorted = idx_unsorted; dp.depth = depth; dp.w0 = w0; dp.w1 = w1;
  dp.face_idx = face_idx; dp.weights = weights; dp.feat = fin; dp.out = fout;
  using K = typename DepthKey<T>::type;
  const size_t smem = (size_t)DEFTET_WARPS * dp.cap * (sizeof(K) + sizeof(unsigned));
  if (smem > 48 * 1024) {
    static std::atomic<int> done[64];                 // once per device (graph-capture safe after warm-up)
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !done[dev].load(std::memory_order_acquire)) {
      CUDA_TRY(cudaFuncSetAttribute(deftet_forward_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    227 * 1024));
      if (dev >= 0 && dev < 64) done[dev].store(1, std::memory_order_release);
    }
  }
  const unsigned nblk = (unsigned)(((int64_t)B * P + DEFTET_WARPS - 1) / DEFTET_WARPS);
  { ProfScope ps(DIBR_KERNEL_DEFTET_FWD, s);
    deftet_forward_kernel<T><<<nblk, DEFTET_WARPS * 32, smem, s>>>(dp, fl, geo, L); }
  LAUNCH_CHECK("deftet_forward_kernel");
  return DIBR_OK;
}

template <typename T>
int deftet_bwd_impl(const T* const* grad_outs, const int64_t* face_idx, const T* weights, const T* img,
                    const T* const* feats, const int* channels, int n_feat, int B, int F, int P, int knum,
                    float eps, T* grad_img, T* const* grad_feats, void* stream) {
  if (B < 0 || F < 0 || P < 0 || knum < 0) return DIBR_ERR_INVALID_ARGUMENT;
  if (n_feat < 0 || n_feat > MAX_FEATS || (n_feat > 0 && (!channels || !feats))) return DIBR_ERR_INVALID_ARGUMENT;
  const size_t nf = (size_t)B * F;
  const int64_t ne = (int64_t)B * P * knum;
  if (nf >= (size_t)1 << 31 || ne >= (int64_t)1 << 40) return DIBR_ERR_TOO_LARGE;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  DeftetBwdParams<T> bp;
  bp.face_idx = face_idx; bp.weights = weights; bp.img = img; bp.B = B; bp.F = F; bp.P = P; bp.knum = knum;
  bp.eps = eps; bp.grad_img = grad_img;
  bp.grad_out.n = bp.feat.n = bp.grad_feat.n = n_feat;
  for (int t = 0; t < MAX_FEATS; ++t) {
    bp.grad_out.ptr[t] = nullptr; bp.feat.ptr[t] = nullptr; bp.grad_feat.ptr[t] = nullptr;
    bp.grad_out.ch[t] = bp.feat.ch[t] = bp.grad_feat.ch[t] = 0;
  }
  for (int t = 0; t < n_feat; ++t) {
    if (channels[t] < 0) return DIBR_ERR_INVALID_ARGUMENT;
    bp.grad_out.ch[t] = bp.feat.ch[t] = bp.grad_feat.ch[t] = channels[t];
    if (channels[t] > 0 && nf > 0 && ne > 0 && (!grad_outs || !grad_outs[t] || !feats[t])) return DIBR_ERR_INVALID_ARGUMENT;
    bp.grad_out.ptr[t] = grad_outs ? grad_outs[t] : nullptr;
    bp.feat.ptr[t] = feats[t];
    bp.grad_feat.ptr[t] = grad_feats ? grad_feats[t] : nullptr;
  }
  if (nf > 0) {
    if (grad_img) CUDA_TRY(cudaMemsetAsync(grad_img, 0, nf * 6 * sizeof(T), s));
    for (int t = 0; t < n_feat; ++t)
      if (channels[t] > 0 && bp.grad_feat.ptr[t])
        CUDA_TRY(cudaMemsetAsync(bp.grad_feat.ptr[t], 0, nf * 3 * channels[t] * sizeof(T), s));
  }
  if (nf == 0 || ne == 0) return DIBR_OK;
  if (!face_idx || !weights || !img) return DIBR_ERR_INVALID_ARGUMENT;
  { ProfScope ps(DIBR_KERNEL_DEFTET_BWD, s);
    deftet_backward_kernel<T><<<(unsigned)((ne + 255) / 256), 256, 0, s>>>(bp); }
  LAUNCH_CHECK("deftet_backward_kernel");
  return DIBR_OK;
}

}  // namespace

#define DIBR_API extern "C" __attribute__((visibility("default")))

DIBR_API int dibr_abi_version(void) { return DIBR_ABI_VERSION; }
DIBR_API const char* dibr_error_string(int code) {
  switch (code) {
    case DIBR_OK: return "ok";
    case DIBR_ERR_INVALID_ARGUMENT: return "invalid argument (null pointer, negative size or unsupported value)";
    case DIBR_ERR_WORKSPACE: return "workspace missing or smaller than dibr_workspace_bytes()";
    case DIBR_ERR_CUDA: return "CUDA error (see dibr_last_cuda_error)";
    case DIBR_ERR_TOO_LARGE: return "dimension too large for the kernels' index types";
    case DIBR_ERR_SCRATCH_LIMIT:
      return "input needs more scratch than the built-in array of the _noalloc entry point holds "
             "(dibr_noalloc_scratch_bytes(); use the variant that takes a caller workspace)";
    default: return "unknown error code";
  }
}
DIBR_API const char* dibr_last_cuda_error(void) { return t_cuda_err; }
DIBR_API uint64_t dibr_kernel_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
DIBR_API void dibr_reset_kernel_launch_count(void) { g_launches.store(0, std::memory_order_relaxed); }
DIBR_API int dibr_profile_next_launch(int kernel_id, void* start_event, void* stop_event) {
  if (kernel_id >= 0 && (!start_event || !stop_event)) return DIBR_ERR_INVALID_ARGUMENT;
  g_prof_start.store(static_cast<cudaEvent_t>(start_event));
  g_prof_stop.store(static_cast<cudaEvent_t>(stop_event));
  g_prof_kernel.store(kernel_id < 0 ? -1 : kernel_id, std::memory_order_release);
  return DIBR_OK;
}

DIBR_API size_t dibr_workspace_bytes(int batch_size, int64_t total_faces, int height, int width) {
  if (batch_size < 0 || total_faces < 0 || height < 0 || width < 0) return 0;
  const Geometry geo = make_geometry(batch_size, height, width);
  const size_t a = ws_layout(batch_size, total_faces, geo).total;                      // rasterizer / deftet layout
  const size_t b = fused_layout(batch_size, total_faces, false, geo).total_no_log;     // soft-mask entry points
  return a > b ? a : b;
}

DIBR_API size_t dibr_soft_mask_workspace_bytes(int batch_size, int64_t total_faces, int height, int width,
                                               int keep_pairs, int elem_size) {
  if (check_dims(batch_size, total_faces, height, width) != DIBR_OK) return 0;
  (void)elem_size;
  const FusedLayout fz = fused_layout(batch_size, total_faces, false, make_geometry(batch_size, height, width));
  return (keep_pairs && fz.log.ok) ? fz.total : fz.total_no_log;
}

DIBR_API size_t dibr_rasterization_workspace_bytes(int batch_size, int64_t total_faces, int height, int width,
                                                   int keep_pairs) {
  if (check_dims(batch_size, total_faces, height, width) != DIBR_OK) return 0;
  const FusedLayout fz = fused_layout(batch_size, total_faces, true, make_geometry(batch_size, height, width));
  return (keep_pairs && fz.log.ok) ? fz.total : fz.total_no_log;
}

DIBR_API size_t dibr_deftet_workspace_bytes(int batch_size, int64_t total_faces) {
  if (batch_size <= 0 || total_faces < 0) return 256;
  const int F = (int)((total_faces + batch_size - 1) / batch_size);
  const int G = deftet_grid(F);
  return ws_layout(batch_size, total_faces, make_geometry(batch_size, G * TILE, G * TILE)).total;
}

DIBR_API size_t dibr_image_l1_workspace_bytes(int64_t n, int elem_size) {
  if (n <= 0 || elem_size <= 0) return 256;
  return align_up((size_t)l1_chunks(n) * (size_t)elem_size);
}
DIBR_API size_t dibr_mask_iou_workspace_bytes(int B, int64_t P, int elem_size) {
  return iou_ws_bytes(B, P, elem_size);
}
DIBR_API size_t dibr_noalloc_scratch_bytes(void) { return NOALLOC_SCRATCH_BYTES; }

#define DEFINE_ENTRY_POINTS(SUF, T)                                                                           \
  DIBR_API int dibr_rasterize_fwd_fused_##SUF(const T* z, const T* img, const T* feat, const uint8_t* valid, \
                                              int B, int F, int C, int H, int W, double multiplier,         \
                                              float eps, T* out_feat, int64_t* out_face_idx,                \
                                              T* out_weights, void* ws, size_t ws_bytes, void* stream) {    \
    if (F < 0 || C < 0) return DIBR_ERR_INVALID_ARGUMENT;                                                     \
    return rasterize_fwd_impl<T>(z, img, nullptr, &feat, &C, 1, valid, nullptr, B, F, (int64_t)B * F, H, W, \
                                 (T)multiplier, 0, (float)multiplier, eps, &out_feat, out_face_idx,          \
                                 out_weights, ws, ws_bytes, stream);                                          \
  }                                                                                                           \
  DIBR_API int dibr_rasterize_fwd_multi_##SUF(const T* z, const T* img, const T* const* feats,               \
                                              const int* channels, int n_feat, const uint8_t* valid, int B, \
                                              int F, int H, int W, double multiplier, float eps,            \
                                              T* const* out_feats, int64_t* out_face_idx, T* out_weights,   \
                                              void* ws, size_t ws_bytes, void* stream) {                     \
    if (F < 0) return DIBR_ERR_INVALID_ARGUMENT;                                                              \
    return rasterize_fwd_impl<T>(z, img, nullptr, feats, channels, n_feat, valid, nullptr, B, F,            \
                                 (int64_t)B * F, H, W, (T)multiplier, 0, (float)multiplier, eps, out_feats,  \
                                 out_face_idx, out_weights, ws, ws_bytes, stream);                            \
  }                                                                                                           \
  DIBR_API int dibr_rasterize_backward_multi_##SUF(                                                          \
      const T* const* grad_outs, const int64_t* face_idx, const T* weights, const T* img,                    \
      const T* const* feats, const int* channels, int n_feat, int B, int F, int H, int W, double multiplier, \
      float eps, T* grad_img, T* const* grad_feats, void* stream) {                                          \
    return rasterize_bwd_impl<T>(grad_outs, face_idx, weights, img, feats, channels, n_feat, B, F, H, W,     \
                                 multiplier, eps, grad_img, grad_feats, stream);                              \
  }                                                                                                           \
  DIBR_API int dibr_packed_rasterize_forward_##SUF(                                                          \
      int H, int W, const T* z, const T* img_mult, const T* bboxes, const T* feat, const int64_t* first_idx, \
      int B, int64_t N, int C, float multiplier, float eps, T* out_feat, int64_t* out_selected_face_idx,     \
      T* out_weights, void* ws, size_t ws_bytes, void* stream) {                                             \
    if (!first_idx || (N > 0 && !bboxes) || C < 0) return DIBR_ERR_INVALID_ARGUMENT;                         \
    return rasterize_fwd_impl<T>(z, img_mult, bboxes, &feat, &C, 1, nullptr, first_idx, B, 0, N, H, W, (T)1, \
                                 1, multiplier, eps, &out_feat, out_selected_face_idx, out_weights, ws,      \
                                 ws_bytes, stream);                                                           \
  }                                                                                                           \
  DIBR_API int dibr_rasterize_backward_##SUF(const T* grad_out, const int64_t* face_idx, const T* weights,   \
                                             const T* img, const T* feat, int B, int F, int C, int H,       \
                                             int W, double multiplier, float eps, T* grad_img,              \
                                             T* grad_feat, void* stream) {                                   \
    if (C < 0) return DIBR_ERR_INVALID_ARGUMENT;                                                              \
    return rasterize_bwd_impl<T>(&grad_out, face_idx, weights, img, &feat, &C, 1, B, F, H, W, multiplier,    \
                                 eps, grad_img, &grad_feat, stream);                                          \
  }                                                                                                           \
  DIBR_API int dibr_soft_mask_fwd_fused_##SUF(const T* img, const int64_t* face_idx, int B, int F, int H,    \
                                              int W, float sigmainv, double boxlen, int knum,               \
                                              double multiplier, T* soft, T* prob, int64_t* close_idx,      \
                                              uint8_t* dist_type, void* ws, size_t ws_bytes,                \
                                              int keep_pairs, void* stream) {                                \
    return soft_mask_impl<T, MODE_FWD>(img, nullptr, face_idx, B, F, H, W, sigmainv, knum, (T)multiplier, 0, \
                                       (T)(boxlen * multiplier), (float)multiplier, soft, prob, close_idx,   \
                                       dist_type, nullptr, nullptr, nullptr, ws, ws_bytes, stream, 0,         \
                                       keep_pairs);                                                           \
  }                                                                                                           \
  DIBR_API int dibr_soft_mask_forward_##SUF(const T* img_mult, const T* large_bboxes,                        \
                                            const int64_t* face_idx, int B, int F, int H, int W,            \
                                            float sigmainv, int knum, float multiplier, T* soft, T* prob,   \
                                            int64_t* close_idx, uint8_t* dist_type, void* ws,               \
                                            size_t ws_bytes, void* stream) {                                 \
    if ((int64_t)B * F > 0 && !large_bboxes) return DIBR_ERR_INVALID_ARGUMENT;                               \
    return soft_mask_impl<T, MODE_FWD>(img_mult, large_bboxes, face_idx, B, F, H, W, sigmainv, knum, (T)1,  \
                                       1, (T)0, multiplier, soft, prob, close_idx, dist_type, nullptr,      \
                                       nullptr, nullptr, ws, ws_bytes, stream);                               \
  }                                                                                                           \
  DIBR_API int dibr_soft_mask_backward_##SUF(const T* grad_soft, const T* soft, const int64_t* face_idx,     \
                                             const T* prob, const int64_t* close_idx,                        \
                                             const uint8_t* dist_type, const T* img_mult, int B, int F,     \
                                             int H, int W, int knum, float sigmainv, float multiplier,      \
                                             T* grad_img, void* stream) {                                    \
    return soft_mask_bwd_sidebuf_impl<T>(grad_soft, soft, face_idx, prob, close_idx, dist_type, img_mult, B, \
                                         F, H, W, knum, sigmainv, multiplier, grad_img, stream);             \
  }                                                                                                           \
  DIBR_API int dibr_soft_mask_bwd_fused_##SUF(const T* grad_soft, const T* soft, const int64_t* face_idx,    \
                                              const T* img, int B, int F, int H, int W, float sigmainv,     \
                                              double boxlen, int knum, double multiplier, T* grad_img,      \
                                              void* ws, size_t ws_bytes, int bins_ready, void* stream) {     \
    return soft_mask_impl<T, MODE_BWD>(img, nullptr, face_idx, B, F, H, W, sigmainv, knum, (T)multiplier, 0, \
                                       (T)(boxlen * multiplier), (float)multiplier, nullptr, nullptr,        \
                                       nullptr, nullptr, grad_soft, soft, grad_img, ws, ws_bytes, stream,    \
                                       bins_ready);                                                           \
  }                                                                                                           \
  DIBR_API int dibr_rasterization_fwd_fused_##SUF(                                                            \
      const T* z, const T* img, const T* const* feats, const int* channels, int n_feat, const T* normals_z,   \
      const uint8_t* valid, int B, int F, int H, int W, double multiplier, float eps, float sigmainv,         \
      double boxlen, int knum, T* const* out_feats, int64_t* out_face_idx, T* soft, void* ws, size_t ws_bytes, \
      int keep_pairs, void* stream) {                                                                          \
    return dibr_rasterization_fwd_impl<T>(z, img, feats, channels, n_feat, normals_z, valid, B, F, H, W,       \
                                          multiplier, eps, sigmainv, boxlen, knum, out_feats, out_face_idx,    \
                                          soft, ws, ws_bytes, keep_pairs, stream);                              \
  }                                                                                                            \
  DIBR_API int dibr_rasterization_bwd_fused_##SUF(                                                            \
      const T* const* grad_outs, const T* grad_soft, const T* soft, const int64_t* face_idx, const T* img,    \
      const T* const* feats, const int* channels, int n_feat, int B, int F, int H, int W, double multiplier,  \
      float eps, float sigmainv, double boxlen, int knum, T* grad_img, T* const* grad_feats, void* ws,        \
      size_t ws_bytes, int bins_ready, void* stream) {                                                         \
    return dibr_rasterization_bwd_impl<T>(grad_outs, grad_soft, soft, face_idx, img, feats, channels, n_feat, \
                                          B, F, H, W, multiplier, eps, sigmainv, boxlen, knum, grad_img,       \
                                          grad_feats, ws, ws_bytes, bins_ready, stream);                        \
  }                                                                                                            \
  DIBR_API int dibr_rasterization_from_mesh_fwd_##SUF(                                                        \
      const T* vertices, const int64_t* faces, const T* camera_proj, const T* rot, const T* trans,            \
      const T* transform, int Bv, int V, const T* const* feats, const int* channels, int n_feat, int B, int F, \
      int H, int W, double multiplier, float eps, float sigmainv, double boxlen, int knum,                     \
      T* face_vertices_image, T* face_vertices_z, T* const* out_feats, int64_t* out_face_idx, T* soft,         \
      void* ws, size_t ws_bytes, int keep_pairs, void* stream) {                                               \
    return from_mesh_fwd_impl<T>(vertices, faces, camera_proj, rot, trans, transform, Bv, V, feats, channels,  \
                                 n_feat, B, F, H, W, multiplier, eps, sigmainv, boxlen, knum,                  \
                                 face_vertices_image, face_vertices_z, out_feats, out_face_idx, soft, ws,      \
                                 ws_bytes, keep_pairs, stream);                                                \
  }                                                                                                            \
  DIBR_API int dibr_rasterization_from_mesh_bwd_##SUF(                                                        \
      const T* const* grad_outs, const T* grad_soft, const T* soft, const int64_t* face_idx,                   \
      const T* face_vertices_image, const T* const* feats, const int* channels, int n_feat,                    \
      const T* vertices, const int64_t* faces, const T* camera_proj, const T* rot, const T* trans,             \
      const T* transform, int Bv, int V, int B, int F, int H, int W, double multiplier, float eps,             \
      float sigmainv, double boxlen, int knum, T* grad_image_scratch, T* const* grad_feats,                    \
      T* grad_vertices, T* grad_proj, T* grad_rot, T* grad_trans, T* grad_transform, void* ws,                 \
      size_t ws_bytes, int bins_ready, void* stream) {                                                         \
    return from_mesh_bwd_impl<T>(grad_outs, grad_soft, soft, face_idx, face_vertices_image, feats, channels,   \
                                 n_feat, vertices, faces, camera_proj, rot, trans, transform, Bv, V, B, F, H,  \
                                 W, multiplier, eps, sigmainv, boxlen, knum, grad_image_scratch, grad_feats,   \
                                 grad_vertices, grad_proj, grad_rot, grad_trans, grad_transform, ws, ws_bytes, \
                                 bins_ready, stream);                                                          \
  }                                                                                                            \
  DIBR_API int dibr_prepare_vertices_fwd_##SUF(const T* vertices, const int64_t* faces, const T* camera_proj, \
                                               const T* rot, const T* trans, const T* transform, int B,      \
                                               int Bv, int V, int F, T* face_vertices_camera,                \
                                               T* face_vertices_image, T* face_normals, void* stream) {      \
    return prepare_impl<T>(false, nullptr, nullptr, nullptr, vertices, faces, camera_proj, rot, trans,       \
                           transform, B, Bv, V, F, face_vertices_camera, face_vertices_image, face_normals,  \
                           nullptr, nullptr, nullptr, nullptr, nullptr, stream);                              \
  }                                                                                                           \
  DIBR_API int dibr_prepare_vertices_bwd_##SUF(const T* grad_camera, const T* grad_image,                    \
                                               const T* grad_normals, const T* vertices,                     \
                                               const int64_t* faces, const T* camera_proj, const T* rot,     \
                                               const T* trans, const T* transform, int B, int Bv, int V,     \
                                               int F, T* grad_vertices, T* grad_proj, T* grad_rot,           \
                                               T* grad_trans, T* grad_transform, void* stream) {             \
    return prepare_impl<T>(true, grad_camera, grad_image, grad_normals, vertices, faces, camera_proj, rot,   \
                           trans, transform, B, Bv, V, F, nullptr, nullptr, nullptr, grad_vertices,          \
                           grad_proj, grad_rot, grad_trans, grad_transform, stream);                          \
  }                                                                                                           \
  DIBR_API int dibr_texture_mapping_fwd_##SUF(const T* uv, const T* texture_maps, const T* mask, int B,      \
                                              int64_t N, int Bt, int C, int Ht, int Wt, int mode, T* out,    \
                                              void* stream) {                                                \
    return texture_impl<T>(false, nullptr, uv, texture_maps, mask, B, N, Bt, C, Ht, Wt, mode, out, nullptr,  \
                           nullptr, nullptr, stream);                                                         \
  }                                                                                                           \
  DIBR_API int dibr_texture_mapping_bwd_##SUF(const T* grad_out, const T* uv, const T* texture_maps,         \
                                              const T* mask, int B, int64_t N, int Bt, int C, int Ht,        \
                                              int Wt, int mode, T* grad_uv, T* grad_texture_maps,            \
                                              T* grad_mask, void* stream) {                                  \
    return texture_impl<T>(true, grad_out, uv, texture_maps, mask, B, N, Bt, C, Ht, Wt, mode, nullptr,       \
                           grad_uv, grad_texture_maps, grad_mask, stream);                                    \
  }                                                                                                           \
  DIBR_API int dibr_adam_step_##SUF(T* param, const T* grad, T* exp_avg, T* exp_avg_sq, float* step,           \
                                    unsigned* arrivals, int64_t n, double lr, double beta1, double beta2,      \
                                    double eps,                                                                \
                                    void* stream) {                                                            \
    return adam_impl<T>(param, grad, exp_avg, exp_avg_sq, step, arrivals, n, lr, beta1, beta2, eps, stream);   \
  }                                                                                                           \
  DIBR_API int dibr_image_l1_fwd_##SUF(const T* a, const T* b, int64_t n, T* loss, void* ws, size_t ws_bytes, \
                                       void* stream) {                                                       \
    return l1_fwd_impl<T>(a, b, n, loss, ws, ws_bytes, stream);                                               \
  }                                                                                                           \
  DIBR_API int dibr_image_l1_bwd_##SUF(const T* grad_loss, const T* a, const T* b, int64_t n, T* grad_a,      \
                                       T* grad_b, void* stream) {                                            \
    return l1_bwd_impl<T>(grad_loss, a, b, n, grad_a, grad_b, stream);                                        \
  }                                                                                                           \
  DIBR_API int dibr_mask_iou_fwd_##SUF(const T* lhs, const T* rhs, int B, int64_t P, T* loss, T* sums,       \
                                       void* ws, size_t ws_bytes, void* stream) {                            \
    return iou_fwd_impl<T>(lhs, rhs, B, P, loss, sums, ws, ws_bytes, stream);                                 \
  }                                                                                                           \
  DIBR_API int dibr_mask_iou_fwd_noalloc_##SUF(const T* lhs, const T* rhs, int B, int64_t P, T* loss,        \
                                               T* sums, void* stream) {                                      \
    return iou_fwd_noalloc_impl<T>(lhs, rhs, B, P, loss, sums, stream);                                       \
  }                                                                                                           \
  DIBR_API int dibr_mask_iou_bwd_##SUF(const T* grad_loss, const T* lhs, const T* rhs, const T* sums,        \
                                       int B, int64_t P, T* grad_lhs, T* grad_rhs, void* stream) {           \
    return iou_bwd_impl<T>(grad_loss, lhs, rhs, sums, B, P, grad_lhs, grad_rhs, stream);                      \
  }                                                                                                           \
  DIBR_API int dibr_deftet_forward_##SUF(const T* z, const T* img, const T* bboxes, const T* pixel_coords,   \
                                         const T* render_ranges, int B, int F, int P, int knum, float eps,   \
                                         int64_t* face_idx, T* pixel_depth, T* w0, T* w1, void* ws,          \
                                         size_t ws_bytes, void* stream) {                                    \
    if ((int64_t)B * F > 0 && !bboxes) return DIBR_ERR_INVALID_ARGUMENT;                                      \
    if ((int64_t)B * P * knum > 0 && !face_idx) return DIBR_ERR_INVALID_ARGUMENT;                             \
    return deftet_fwd_impl<T>(z, img, bboxes, pixel_coords, render_ranges, nullptr, nullptr, 0, B, F, P,     \
                              knum, eps, nullptr, nullptr, nullptr, face_idx, pixel_depth, w0, w1, ws,       \
                              ws_bytes, stream);                                                              \
  }                                                                                                           \
  DIBR_API int dibr_deftet_render_fused_##SUF(const T* pixel_coords, const T* render_ranges, const T* z,     \
                                              const T* img, const T* const* feats, const int* channels,      \
                                              int n_feat, int B, int F, int P, int knum, float eps,          \
                                              T* const* outs, int64_t* face_idx, T* weights, void* ws,       \
                                              size_t ws_bytes, void* stream) {                               \
    return deftet_fwd_impl<T>(z, img, nullptr, pixel_coords, render_ranges, feats, channels, n_feat, B, F,   \
                              P, knum, eps, outs, face_idx, weights, nullptr, nullptr, nullptr, nullptr, ws, \
                              ws_bytes, stream);                                                              \
  }                                                                                                           \
  DIBR_API int dibr_deftet_backward_##SUF(const T* const* grad_outs, const int64_t* face_idx,                \
                                          const T* weights, const T* img, const T* const* feats,             \
                                          const int* channels, int n_feat, int B, int F, int P, int knum,    \
                                          float eps, T* grad_img, T* const* grad_feats, void* stream) {      \
    return deftet_bwd_impl<T>(grad_outs, face_idx, weights, img, feats, channels, n_feat, B, F, P, knum,     \
                              eps, grad_img, grad_feats, stream);                                             \
  }

DEFINE_ENTRY_POINTS(f32, float)
DEFINE_ENTRY_POINTS(f64, double)
