"""Loader / builder of libdibr_b200.so (the C-ABI declared in include/dibr_b200.h).

The product path has NO fallback: if the CUDA library is missing or fails to load, every op raises.
PyTorch is used here only as plumbing (device memory, current stream); the library itself does not
link against torch.
"""
import ctypes
import os
import subprocess
import threading

import torch  # noqa: F401  (imported first so that libcudart.so.12 is already mapped)

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
# DIBR_B200_LIB selects another build of the same library (kernel-tuning experiments); the default is in-tree
LIB_PATH = os.environ.get("DIBR_B200_LIB") or os.path.join(_PKG, "libdibr_b200.so")
CSRC = os.path.join(_PKG, "csrc")
SOURCES = ["dibr_c_abi.cu"]
HEADERS = ["dibr_common.cuh", "dibr_binning.cuh", "dibr_raster_fwd.cuh", "dibr_raster_bwd.cuh",
           "dibr_soft_mask.cuh", "dibr_prepare.cuh", "dibr_texture.cuh", "dibr_losses.cuh", "dibr_deftet.cuh"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-cudart", "shared", "-Xcompiler", "-fPIC,-fvisibility=hidden", "-ldl"]

_lock = threading.Lock()
_lib = None

c_f, c_d, c_i, c_i64, c_p, c_sz = (ctypes.c_float, ctypes.c_double, ctypes.c_int, ctypes.c_int64,
                                   ctypes.c_void_p, ctypes.c_size_t)

# name (without _f32/_f64 suffix) -> argtypes; mirrors include/dibr_b200.h
_TYPED = {
    "dibr_rasterize_fwd_fused": [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_d, c_f, c_p, c_p, c_p, c_p, c_sz, c_p],
    "dibr_packed_rasterize_forward": [c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_i64, c_i, c_f, c_f, c_p, c_p, c_p,
                                      c_p, c_sz, c_p],
    "dibr_rasterize_backward": [c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_d, c_f, c_p, c_p, c_p],
    "dibr_rasterize_fwd_multi": [c_p, c_p, c_p, c_p, c_i, c_p, c_i, c_i, c_i, c_i, c_d, c_f, c_p, c_p, c_p, c_p, c_sz, c_p],
    "dibr_rasterize_backward_multi": [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_d, c_f, c_p, c_p, c_p],
    "dibr_soft_mask_fwd_fused": [c_p, c_p, c_i, c_i, c_i, c_i, c_f, c_d, c_i, c_d, c_p, c_p, c_p, c_p, c_p, c_sz, c_i,
                                 c_p],
    "dibr_soft_mask_forward": [c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_f, c_i, c_f, c_p, c_p, c_p, c_p, c_p, c_sz, c_p],
    "dibr_soft_mask_backward": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_f, c_f, c_p, c_p],
    "dibr_soft_mask_bwd_fused": [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_f, c_d, c_i, c_d, c_p, c_p, c_sz, c_i, c_p],
    "dibr_rasterization_fwd_fused": [c_p, c_p, c_p, c_p, c_i, c_p, c_p, c_i, c_i, c_i, c_i, c_d, c_f, c_f, c_d, c_i, c_p,
                                     c_p, c_p, c_p, c_sz, c_i, c_p],
    "dibr_rasterization_bwd_fused": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_d, c_f, c_f, c_d,
                                     c_i, c_p, c_p, c_p, c_sz, c_i, c_p],
    "dibr_rasterization_from_mesh_fwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_d,
                                         c_f, c_f, c_d, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_sz, c_i, c_p],
    "dibr_rasterization_from_mesh_bwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i,
                                         c_i, c_i, c_i, c_i, c_d, c_f, c_f, c_d, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p,
                                         c_p, c_sz, c_i, c_p],
    "dibr_prepare_vertices_fwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_p, c_p, c_p, c_p],
    "dibr_prepare_vertices_bwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_p, c_p, c_p, c_p,
                                  c_p, c_p],
    "dibr_texture_mapping_fwd": [c_p, c_p, c_p, c_i, c_i64, c_i, c_i, c_i, c_i, c_i, c_p, c_p],
    "dibr_texture_mapping_bwd": [c_p, c_p, c_p, c_p, c_i, c_i64, c_i, c_i, c_i, c_i, c_i, c_p, c_p, c_p, c_p],
    "dibr_adam_step": [c_p, c_p, c_p, c_p, c_p, c_p, c_i64, c_d, c_d, c_d, c_d, c_p],
    "dibr_image_l1_fwd": [c_p, c_p, c_i64, c_p, c_p, c_sz, c_p],
    "dibr_image_l1_bwd": [c_p, c_p, c_p, c_i64, c_p, c_p, c_p],
    "dibr_mask_iou_fwd": [c_p, c_p, c_i, c_i64, c_p, c_p, c_p, c_sz, c_p],
    "dibr_mask_iou_fwd_noalloc": [c_p, c_p, c_i, c_i64, c_p, c_p, c_p],
    "dibr_mask_iou_bwd": [c_p, c_p, c_p, c_p, c_i, c_i64, c_p, c_p, c_p],
    "dibr_deftet_forward": [c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_f, c_p, c_p, c_p, c_p, c_p, c_sz, c_p],
    "dibr_deftet_render_fused": [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_f, c_p, c_p, c_p, c_p,
                                 c_sz, c_p],
    "dibr_deftet_backward": [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_f, c_p, c_p, c_p],
}
EXPORTED_SYMBOLS = (["dibr_abi_version", "dibr_error_string", "dibr_last_cuda_error", "dibr_workspace_bytes",
                     "dibr_kernel_launch_count", "dibr_reset_kernel_launch_count", "dibr_profile_next_launch",
                     "dibr_mask_iou_workspace_bytes", "dibr_image_l1_workspace_bytes", "dibr_noalloc_scratch_bytes", "dibr_deftet_workspace_bytes",
                     "dibr_soft_mask_workspace_bytes", "dibr_rasterization_workspace_bytes"] +
                    [f"{n}_{s}" for n in _TYPED for s in ("f32", "f64")])


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.join(_ROOT, "include", "dibr_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    """Compile the CUDA sources for sm_100a into kaolin_b200/libdibr_b200.so (in-tree)."""
    if not force and not needs_build():
        return LIB_PATH
    cmd = ["nvcc", "-shared"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB_PATH] + [os.path.join(CSRC, f) for f in SOURCES]
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or out.returncode != 0:
        print(out.stdout)
    if out.returncode != 0:
        raise RuntimeError("nvcc failed building libdibr_b200.so:\n" + out.stdout)
    return LIB_PATH


def load():
    """ctypes handle of libdibr_b200.so; raises if it is absent (never falls back to another path)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"kaolin_b200: {LIB_PATH} not found — the CUDA extension is required (no CPU/PyTorch "
                "fallback exists). Build it with `python -c 'import __graft_entry__ as g; g.build()'`.")
        lib = ctypes.CDLL(LIB_PATH)
        lib.dibr_abi_version.restype = c_i
        lib.dibr_error_string.restype = ctypes.c_char_p
        lib.dibr_error_string.argtypes = [c_i]
        lib.dibr_last_cuda_error.restype = ctypes.c_char_p
        lib.dibr_workspace_bytes.restype = c_sz
        lib.dibr_workspace_bytes.argtypes = [c_i, c_i64, c_i, c_i]
        lib.dibr_kernel_launch_count.restype = ctypes.c_uint64
        lib.dibr_reset_kernel_launch_count.restype = None
        lib.dibr_mask_iou_workspace_bytes.restype = c_sz
        lib.dibr_mask_iou_workspace_bytes.argtypes = [c_i, c_i64, c_i]
        lib.dibr_image_l1_workspace_bytes.restype = c_sz
        lib.dibr_image_l1_workspace_bytes.argtypes = [c_i64, c_i]
        lib.dibr_soft_mask_workspace_bytes.restype = c_sz
        lib.dibr_soft_mask_workspace_bytes.argtypes = [c_i, c_i64, c_i, c_i, c_i, c_i]
        lib.dibr_noalloc_scratch_bytes.restype = c_sz
        lib.dibr_noalloc_scratch_bytes.argtypes = []
        lib.dibr_rasterization_workspace_bytes.restype = c_sz
        lib.dibr_rasterization_workspace_bytes.argtypes = [c_i, c_i64, c_i, c_i, c_i]
        lib.dibr_deftet_workspace_bytes.restype = c_sz
        lib.dibr_deftet_workspace_bytes.argtypes = [c_i, c_i64]
        lib.dibr_profile_next_launch.restype = c_i
        lib.dibr_profile_next_launch.argtypes = [c_i, c_p, c_p]
        for name, argtypes in _TYPED.items():
            for suf in ("f32", "f64"):
                fn = getattr(lib, f"{name}_{suf}")
                fn.restype = c_i
                fn.argtypes = argtypes
        if lib.dibr_abi_version() != 8:
            raise RuntimeError("kaolin_b200: libdibr_b200.so ABI version mismatch")
        _lib = lib
    return _lib


def check(code, func_name):
    """Map a non-zero return code to the reference's error behaviour (RuntimeError)."""
    if code != 0:
        lib = load()
        msg = lib.dibr_error_string(code).decode()
        if code == 3:
            msg += ": " + lib.dibr_last_cuda_error().decode()
        raise RuntimeError(f"{func_name}: {msg}")


def suffix(dtype, func_name):
    if dtype == torch.float32:
        return "f32"
    if dtype == torch.float64:
        return "f64"
    # reference: AT_ERROR("<name> not implemented for '<type>'"), rasterization_cuda.cu:36-37
    raise RuntimeError(f"{func_name} not implemented for '{str(dtype).replace('torch.', '').capitalize()}'")


def ptr(t):
    return 0 if t is None else t.data_ptr()


def ptr_array(tensors):
    """HOST array of device pointers (for the *_multi entry points)."""
    return (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def int_array(values):
    return (ctypes.c_int * len(values))(*[int(v) for v in values])


def workspace(batch_size, total_faces, height, width, device):
    n = load().dibr_workspace_bytes(int(batch_size), int(total_faces), int(height), int(width))
    return torch.empty(max(int(n), 256), dtype=torch.uint8, device=device), n


def current_stream(device):
    return torch.cuda.current_stream(device).cuda_stream


KERNEL_IDS = dict(setup_count=0, scan=1, fill=2, raster_fwd=3, raster_bwd=4, soft_fwd=5, soft_bwd=6,
                  prepare_fwd=7, prepare_bwd=8, texture_fwd=9, texture_bwd=10, iou_fwd=11, iou_bwd=12,
                  deftet_fwd=13, deftet_bwd=14)


def profile_next_launch(kernel_id, start_event, stop_event):
    """Arm the per-kernel timing hook with two torch.cuda.Event objects (created with enable_timing=True and
    already recorded once so that their handles exist)."""
    code = load().dibr_profile_next_launch(int(kernel_id), start_event.cuda_event if start_event else None,
                                           stop_event.cuda_event if stop_event else None)
    check(code, "dibr_profile_next_launch")


def kernel_launch_count():
    return int(load().dibr_kernel_launch_count())


def reset_kernel_launch_count():
    load().dibr_reset_kernel_launch_count()
