"""``kaolin_b200._C.render.mesh`` — the reference's native ops of the path (kaolin/csrc/bindings.cpp:75-82),
same names, positional arguments, return conventions and error behaviour, backed by libdibr_b200.so.

Validation mirrors the host wrappers kaolin/csrc/render/mesh/rasterization.cpp:60-85,127-154 and
dibr_soft_mask.cpp:56-82,135-170 (at::checkAllSameGPU / checkAllContiguous / checkSize -> RuntimeError).
Outputs are freshly allocated with the input's options; inputs are never mutated; no state is kept.
"""
import types

import torch

from . import _lib


def _check_all_same_gpu(fn, named):
    dev = None
    for name, t in named:
        if not isinstance(t, torch.Tensor):
            raise TypeError(f"{fn}(): argument '{name}' must be a Tensor")
        if not t.is_cuda:
            raise RuntimeError(f"Expected tensor for argument '{name}' to have CUDA type; but got {t.device.type} "
                               f"instead (while checking arguments for {fn})")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"Expected tensor for argument '{name}' to be on GPU {dev.index} (while checking "
                               f"arguments for {fn})")
    return dev


def _check_all_contiguous(fn, named):
    for name, t in named:
        if not t.is_contiguous():
            raise RuntimeError(f"Expected contiguous tensor, but got non-contiguous tensor for argument '{name}' "
                               f"(while checking arguments for {fn})")


def _check_size(fn, name, t, sizes):
    if tuple(t.shape) != tuple(sizes):
        raise RuntimeError(f"Expected tensor of size {list(sizes)}, but got tensor of size {list(t.shape)} for "
                           f"argument '{name}' (while checking arguments for {fn})")


def _check_dtype(fn, name, t, dtype):
    if t.dtype != dtype:
        raise RuntimeError(f"{fn}: expected '{name}' to have dtype {dtype} but got {t.dtype}")


def packed_rasterize_forward_cuda(height, width, face_vertices_z, face_vertices_image, face_bboxes, face_features,
                                  first_idx_face_per_mesh, multiplier, eps):
    """rasterization.h:23-32.  Returns [interpolated_features, selected_face_idx, output_weights]."""
    fn = "packed_rasterize_forward_cuda"
    named = [("face_vertices_z", face_vertices_z), ("face_vertices_image", face_vertices_image),
             ("face_bboxes", face_bboxes), ("face_features", face_features),
             ("first_idx_face_per_mesh", first_idx_face_per_mesh)]
    dev = _check_all_same_gpu(fn, named)
    _check_all_contiguous(fn, named)
    n = face_vertices_z.shape[0]
    b = first_idx_face_per_mesh.shape[0] - 1
    c = face_features.shape[2] if face_features.dim() == 3 else -1
    _check_size(fn, "face_vertices_z", face_vertices_z, (n, 3))
    _check_size(fn, "face_vertices_image", face_vertices_image, (n, 3, 2))
    _check_size(fn, "face_bboxes", face_bboxes, (n, 4))
    _check_size(fn, "face_features", face_features, (n, 3, c))
    _check_size(fn, "first_idx_face_per_mesh", first_idx_face_per_mesh, (b + 1,))
    suf = _lib.suffix(face_vertices_z.dtype, fn)
    for nm, t in named[1:4]:
        _check_dtype(fn, nm, t, face_vertices_z.dtype)
    _check_dtype(fn, "first_idx_face_per_mesh", first_idx_face_per_mesh, torch.long)
    opt = dict(dtype=face_vertices_z.dtype, device=dev)
    sel = torch.empty((b, height, width), dtype=torch.long, device=dev)
    w = torch.empty((b, height, width, 3), **opt)
    feat = torch.empty((b, height, width, c), **opt)
    with torch.cuda.device(dev):
        ws, nbytes = _lib.workspace(b, n, height, width, dev)
        code = getattr(_lib.load(), "dibr_packed_rasterize_forward_" + suf)(
            int(height), int(width), _lib.ptr(face_vertices_z), _lib.ptr(face_vertices_image), _lib.ptr(face_bboxes),
            _lib.ptr(face_features), _lib.ptr(first_idx_face_per_mesh), b, n, c, float(multiplier), float(eps),
            _lib.ptr(feat), _lib.ptr(sel), _lib.ptr(w), _lib.ptr(ws), nbytes, _lib.current_stream(dev))
    _lib.check(code, fn)
    return [feat, sel, w]


def rasterize_backward_cuda(grad_interpolated_features, interpolated_features, selected_face_idx, output_weights,
                            face_vertices_image, face_features, eps):
    """rasterization.h:34-41.  Returns [grad_face_vertices_image, grad_face_features]."""
    fn = "rasterize_backward_cuda"
    named = [("grad_interpolated_features", grad_interpolated_features),
             ("interpolated_features", interpolated_features), ("selected_face_idx", selected_face_idx),
             ("output_weights", output_weights), ("face_vertices_image", face_vertices_image),
             ("face_features", face_features)]
    dev = _check_all_same_gpu(fn, named)
    _check_all_contiguous(fn, named)
    b, h, w, c = grad_interpolated_features.shape
    f = face_vertices_image.shape[1]
    _check_size(fn, "interpolated_features", interpolated_features, (b, h, w, c))
    _check_size(fn, "selected_face_idx", selected_face_idx, (b, h, w))
    _check_size(fn, "output_weights", output_weights, (b, h, w, 3))
    _check_size(fn, "face_vertices_image", face_vertices_image, (b, f, 3, 2))
    _check_size(fn, "face_features", face_features, (b, f, 3, c))
    suf = _lib.suffix(grad_interpolated_features.dtype, fn)
    for nm, t in (named[3], named[4], named[5]):
        _check_dtype(fn, nm, t, grad_interpolated_features.dtype)
    _check_dtype(fn, "selected_face_idx", selected_face_idx, torch.long)
    gi = torch.empty_like(face_vertices_image)
    gf = torch.empty_like(face_features)
    with torch.cuda.device(dev):
        code = getattr(_lib.load(), "dibr_rasterize_backward_" + suf)(
            _lib.ptr(grad_interpolated_features), _lib.ptr(selected_face_idx), _lib.ptr(output_weights),
            _lib.ptr(face_vertices_image), _lib.ptr(face_features), b, f, c, h, w, 1.0, float(eps),
            _lib.ptr(gi), _lib.ptr(gf), _lib.current_stream(dev))
    _lib.check(code, fn)
    return [gi, gf]


def dibr_soft_mask_forward_cuda(face_vertices_image, face_large_bboxes, selected_face_idx, sigmainv, knum,
                                multiplier):
    """dibr_soft_mask.h:23-29.  Returns [soft_mask, close_face_prob, close_face_idx, close_face_dist_type]."""
    fn = "dibr_soft_mask_forward_cuda"
    named = [("face_vertices_image", face_vertices_image), ("face_bboxes", face_large_bboxes),
             ("selected_face_idx", selected_face_idx)]
    dev = _check_all_same_gpu(fn, named)
    _check_all_contiguous(fn, named)
    b, f = face_vertices_image.shape[0], face_vertices_image.shape[1]
    h, w = selected_face_idx.shape[1], selected_face_idx.shape[2]
    _check_size(fn, "face_vertices_image", face_vertices_image, (b, f, 3, 2))
    _check_size(fn, "face_bboxes", face_large_bboxes, (b, f, 4))
    _check_size(fn, "selected_face_idx", selected_face_idx, (b, h, w))
    suf = _lib.suffix(face_vertices_image.dtype, fn)
    _check_dtype(fn, "face_bboxes", face_large_bboxes, face_vertices_image.dtype)
    _check_dtype(fn, "selected_face_idx", selected_face_idx, torch.long)
    knum = int(knum)
    opt = dict(dtype=face_vertices_image.dtype, device=dev)
    cidx = torch.empty((b, h, w, knum), dtype=torch.long, device=dev)
    prob = torch.empty((b, h, w, knum), **opt)
    ctype = torch.empty((b, h, w, knum), dtype=torch.uint8, device=dev)
    soft = torch.empty((b, h, w), **opt)
    with torch.cuda.device(dev):
        ws, nbytes = _lib.workspace(b, b * f, h, w, dev)
        code = getattr(_lib.load(), "dibr_soft_mask_forward_" + suf)(
            _lib.ptr(face_vertices_image), _lib.ptr(face_large_bboxes), _lib.ptr(selected_face_idx), b, f, h, w,
            float(sigmainv), knum, float(multiplier), _lib.ptr(soft), _lib.ptr(prob), _lib.ptr(cidx),
            _lib.ptr(ctype), _lib.ptr(ws), nbytes, _lib.current_stream(dev))
    _lib.check(code, fn)
    return [soft, prob, cidx, ctype]


def dibr_soft_mask_backward_cuda(grad_soft_mask, soft_mask, selected_face_idx, close_face_prob, close_face_idx,
                                 close_face_dist_type, face_vertices_image, sigmainv, multiplier):
    """dibr_soft_mask.h:31-40.  Returns grad_face_vertices_image."""
    fn = "dibr_soft_mask_backward_cuda"
    named = [("grad_soft_mask", grad_soft_mask), ("soft_mask", soft_mask), ("selected_face_idx", selected_face_idx),
             ("close_face_prob", close_face_prob), ("close_face_idx", close_face_idx),
             ("close_face_dist_type", close_face_dist_type), ("face_vertices_image", face_vertices_image)]
    dev = _check_all_same_gpu(fn, named)
    _check_all_contiguous(fn, named)
    b, f = face_vertices_image.shape[0], face_vertices_image.shape[1]
    h, w = selected_face_idx.shape[1], selected_face_idx.shape[2]
    knum = close_face_idx.shape[-1]
    _check_size(fn, "grad_soft_mask", grad_soft_mask, (b, h, w))
    _check_size(fn, "soft_mask", soft_mask, (b, h, w))
    _check_size(fn, "selected_face_idx", selected_face_idx, (b, h, w))
    _check_size(fn, "close_face_prob", close_face_prob, (b, h, w, knum))
    _check_size(fn, "close_face_idx", close_face_idx, (b, h, w, knum))
    _check_size(fn, "close_face_dist_type", close_face_dist_type, (b, h, w, knum))
    _check_size(fn, "face_vertices_image", face_vertices_image, (b, f, 3, 2))
    suf = _lib.suffix(face_vertices_image.dtype, fn)
    for nm, t in (named[0], named[1], named[3]):
        _check_dtype(fn, nm, t, face_vertices_image.dtype)
    _check_dtype(fn, "selected_face_idx", selected_face_idx, torch.long)
    _check_dtype(fn, "close_face_idx", close_face_idx, torch.long)
    _check_dtype(fn, "close_face_dist_type", close_face_dist_type, torch.uint8)
    gi = torch.empty_like(face_vertices_image)
    with torch.cuda.device(dev):
        code = getattr(_lib.load(), "dibr_soft_mask_backward_" + suf)(
            _lib.ptr(grad_soft_mask), _lib.ptr(soft_mask), _lib.ptr(selected_face_idx), _lib.ptr(close_face_prob),
            _lib.ptr(close_face_idx), _lib.ptr(close_face_dist_type), _lib.ptr(face_vertices_image), b, f, h, w,
            knum, float(sigmainv), float(multiplier), _lib.ptr(gi), _lib.current_stream(dev))
    _lib.check(code, fn)
    return gi


def deftet_sparse_render_forward_cuda(face_vertices_z, face_vertices_image, face_bboxes, pixel_coords,
                                      pixel_depth_ranges, knum, eps):
    """kaolin/csrc/render/mesh/deftet.h (deftet.cpp:47-109).  Returns [selected_face_idx, pixel_depths, w0, w1]
    in face order, with the reference's fills (-1, -inf, 0, 0)."""
    fn = "deftet_sparse_render_forward_cuda"
    named = [("face_vertices_z", face_vertices_z), ("face_vertices_image", face_vertices_image),
             ("face_bboxes", face_bboxes), ("pixel_coords", pixel_coords),
             ("pixel_depth_ranges", pixel_depth_ranges)]
    dev = _check_all_same_gpu(fn, named)
    _check_all_contiguous(fn, named)
    b, f = face_vertices_z.shape[0], face_vertices_z.shape[1]
    p = pixel_coords.shape[1]
    _check_size(fn, "face_vertices_z", face_vertices_z, (b, f, 3))
    _check_size(fn, "face_vertices_image", face_vertices_image, (b, f, 3, 2))
    _check_size(fn, "face_bboxes", face_bboxes, (b, f, 4))
    _check_size(fn, "pixel_coords", pixel_coords, (b, p, 2))
    _check_size(fn, "pixel_depth_ranges", pixel_depth_ranges, (b, p, 2))
    suf = _lib.suffix(face_vertices_z.dtype, fn)
    for nm, t in named[1:]:
        _check_dtype(fn, nm, t, face_vertices_z.dtype)
    knum = int(knum)
    opt = dict(dtype=face_vertices_z.dtype, device=dev)
    idx = torch.empty((b, p, knum), dtype=torch.long, device=dev)
    depth, w0, w1 = (torch.empty((b, p, knum), **opt) for _ in range(3))
    lib = _lib.load()
    with torch.cuda.device(dev):
        nbytes = lib.dibr_deftet_workspace_bytes(b, b * f)
        ws = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
        code = getattr(lib, "dibr_deftet_forward_" + suf)(
            _lib.ptr(face_vertices_z), _lib.ptr(face_vertices_image), _lib.ptr(face_bboxes), _lib.ptr(pixel_coords),
            _lib.ptr(pixel_depth_ranges), b, f, p, knum, float(eps), _lib.ptr(idx), _lib.ptr(depth), _lib.ptr(w0),
            _lib.ptr(w1), _lib.ptr(ws), nbytes, _lib.current_stream(dev))
    _lib.check(code, fn)
    return [idx, depth, w0, w1]


def deftet_sparse_render_backward_cuda(grad_interpolated_features, face_idx, weights, face_vertices_image,
                                       face_features, eps):
    """kaolin/csrc/render/mesh/deftet.h (deftet.cpp:111-170).  Returns [grad_face_vertices_image,
    grad_face_features]."""
    fn = "deftet_sparse_render_backward_cuda"
    named = [("grad_interpolated_features", grad_interpolated_features), ("face_idx", face_idx),
             ("weights", weights), ("face_vertices_image", face_vertices_image), ("face_features", face_features)]
    dev = _check_all_same_gpu(fn, named)
    _check_all_contiguous(fn, named)
    if grad_interpolated_features.dim() != 4:
        raise RuntimeError(f"{fn}: grad_interpolated_features must be of shape (B,P,knum,C)")
    b, p, knum, c = grad_interpolated_features.shape
    f = face_vertices_image.shape[1]
    _check_size(fn, "face_idx", face_idx, (b, p, knum))
    _check_size(fn, "weights", weights, (b, p, knum, 3))
    _check_size(fn, "face_vertices_image", face_vertices_image, (b, f, 3, 2))
    _check_size(fn, "face_features", face_features, (b, f, 3, c))
    suf = _lib.suffix(face_vertices_image.dtype, fn)
    for nm, t in (named[0], named[2], named[4]):
        _check_dtype(fn, nm, t, face_vertices_image.dtype)
    _check_dtype(fn, "face_idx", face_idx, torch.long)
    gi = torch.empty_like(face_vertices_image)
    gf = torch.empty_like(face_features)
    with torch.cuda.device(dev):
        code = getattr(_lib.load(), "dibr_deftet_backward_" + suf)(
            _lib.ptr_array([grad_interpolated_features]), _lib.ptr(face_idx), _lib.ptr(weights),
            _lib.ptr(face_vertices_image), _lib.ptr_array([face_features]), _lib.int_array([c]), 1, b, f, p, knum,
            float(eps), _lib.ptr(gi), _lib.ptr_array([gf]), _lib.current_stream(dev))
    _lib.check(code, fn)
    return [gi, gf]


mesh = types.SimpleNamespace(
    deftet_sparse_render_forward_cuda=deftet_sparse_render_forward_cuda,
    deftet_sparse_render_backward_cuda=deftet_sparse_render_backward_cuda,
    packed_rasterize_forward_cuda=packed_rasterize_forward_cuda,
    rasterize_backward_cuda=rasterize_backward_cuda,
    dibr_soft_mask_forward_cuda=dibr_soft_mask_forward_cuda,
    dibr_soft_mask_backward_cuda=dibr_soft_mask_backward_cuda,
)
render = types.SimpleNamespace(mesh=mesh)
