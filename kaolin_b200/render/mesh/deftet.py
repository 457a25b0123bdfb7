"""Mirror of ``kaolin.render.mesh.deftet`` (kaolin/render/mesh/deftet.py:269-417): ``deftet_sparse_render`` and
its autograd Function, backed by libdibr_b200.so.

``DeftetSparseRenderer.forward`` of the reference computes bboxes with three torch kernels, runs an O(P*F)
warp-per-point kernel, then argsort + four gathers + a padded feature gather + a weighted sum
(deftet.py:288-312).  Here one C-ABI call: faces are binned on a uniform grid (the rasterizer's two-level
lists), each query point walks only its cell's lists, keeps the first knum faces in face order, sorts them
by depth, and interpolates every feature tensor of a list in place (no ``torch.cat``, no output slicing).
Equal depths keep face order (the reference's ``torch.argsort`` leaves that order unspecified).
"""
import torch
from torch.autograd import Function

from ... import _lib
from .rasterization import _require_cuda

__all__ = ["deftet_sparse_render"]


class DeftetSparseRenderer(Function):
    """forward(ctx, pixel_coords, render_ranges, face_vertices_z, face_vertices_image, knum, eps, *face_features)
    -> (*interpolated_features, sorted_face_idx)     (reference: deftet.py:269-337, one feature tensor there)."""
    @staticmethod
    def forward(ctx, pixel_coords, render_ranges, face_vertices_z, face_vertices_image, knum, eps,
                *face_features):
        fn = "deftet_sparse_render_forward_cuda"
        dev = _require_cuda(fn, pixel_coords=pixel_coords, render_ranges=render_ranges,
                            face_vertices_z=face_vertices_z, face_vertices_image=face_vertices_image,
                            **{f"face_features[{i}]": f for i, f in enumerate(face_features)})
        dtype = face_vertices_z.dtype
        suf = _lib.suffix(dtype, fn)
        if face_vertices_z.dim() != 3 or face_vertices_z.shape[-1] != 3:
            raise RuntimeError(f"{fn}: face_vertices_z must be of shape (B,F,3)")
        b, f = face_vertices_z.shape[0], face_vertices_z.shape[1]
        if pixel_coords.dim() != 3 or pixel_coords.shape[0] != b or pixel_coords.shape[2] != 2:
            raise RuntimeError(f"{fn}: pixel_coords must be of shape (B,P,2)")
        p = pixel_coords.shape[1]
        if tuple(render_ranges.shape) != (b, p, 2) or tuple(face_vertices_image.shape) != (b, f, 3, 2):
            raise RuntimeError(f"{fn}: expected render_ranges (B,P,2) and face_vertices_image (B,F,3,2)")
        if len(face_features) > 8:
            raise RuntimeError(f"{fn}: at most 8 feature tensors per call")
        feats = []
        for t in (pixel_coords, render_ranges, face_vertices_image, *face_features):
            if t.dtype != dtype:
                raise RuntimeError(f"{fn}: all floating point inputs must be of dtype {dtype}")
        for t in face_features:
            if t.dim() != 4 or tuple(t.shape[:3]) != (b, f, 3):
                raise RuntimeError(f"{fn}: every face_features tensor must be of shape (B,F,3,C)")
            feats.append(t.contiguous())
        knum = int(knum)
        pixel_coords, render_ranges = pixel_coords.contiguous(), render_ranges.contiguous()
        face_vertices_z, face_vertices_image = face_vertices_z.contiguous(), face_vertices_image.contiguous()
        chans = [t.shape[-1] for t in feats]
        outs = [torch.empty((b, p, knum, c), dtype=dtype, device=dev) for c in chans]
        face_idx = torch.empty((b, p, knum), dtype=torch.long, device=dev)
        need_bwd = any(ctx.needs_input_grad[3:4]) or any(ctx.needs_input_grad[6:])
        weights = torch.empty((b, p, knum, 3), dtype=dtype, device=dev) if need_bwd else None
        lib = _lib.load()
        with torch.cuda.device(dev):
            nbytes = lib.dibr_deftet_workspace_bytes(b, b * f)
            ws = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
            code = getattr(lib, "dibr_deftet_render_fused_" + suf)(
                _lib.ptr(pixel_coords), _lib.ptr(render_ranges), _lib.ptr(face_vertices_z),
                _lib.ptr(face_vertices_image), _lib.ptr_array(feats), _lib.int_array(chans), len(feats), b, f, p,
                knum, float(eps), _lib.ptr_array(outs), _lib.ptr(face_idx), _lib.ptr(weights), _lib.ptr(ws), nbytes,
                _lib.current_stream(dev))
        _lib.check(code, fn)
        ctx.need_bwd = need_bwd
        if need_bwd:
            ctx.save_for_backward(face_idx, weights, face_vertices_image, *feats)
        ctx.mark_non_differentiable(face_idx)
        ctx.set_materialize_grads(False)
        ctx.eps = eps
        ctx.out_shapes = [tuple(o.shape) for o in outs]
        return (*outs, face_idx)

    @staticmethod
    def backward(ctx, *grads):
        fn = "deftet_sparse_render_backward_cuda"
        n_in = 6 + len(grads) - 1
        # nothing differentiable asked for (e.g. only face_vertices_z / pixel_coords / render_ranges require grad:
        # the reference returns None for those, deftet.py:337): no tensors were saved
        if not ctx.need_bwd or all(g is None for g in grads[:-1]):
            return (None,) * n_in
        face_idx, weights, face_vertices_image = ctx.saved_tensors[:3]
        feats = list(ctx.saved_tensors[3:])
        dev, dtype = face_vertices_image.device, face_vertices_image.dtype
        suf = _lib.suffix(dtype, fn)
        gouts = [torch.zeros(shp, dtype=dtype, device=dev) if g is None else g.contiguous()
                 for g, shp in zip(grads[:-1], ctx.out_shapes)]
        b, f = face_vertices_image.shape[0], face_vertices_image.shape[1]
        p, knum = face_idx.shape[1], face_idx.shape[2]
        gi = torch.empty_like(face_vertices_image)
        gfs = [torch.empty_like(t) for t in feats]
        chans = [t.shape[-1] for t in feats]
        with torch.cuda.device(dev):
            code = getattr(_lib.load(), "dibr_deftet_backward_" + suf)(
                _lib.ptr_array(gouts), _lib.ptr(face_idx), _lib.ptr(weights), _lib.ptr(face_vertices_image),
                _lib.ptr_array(feats), _lib.int_array(chans), len(feats), b, f, p, knum, float(ctx.eps),
                _lib.ptr(gi), _lib.ptr_array(gfs), _lib.current_stream(dev))
        _lib.check(code, fn)
        # reference: (None, None, None, grad_face_vertices_image, grad_face_features, None, None), deftet.py:337
        return (None, None, None, gi, None, None, *gfs)


def deftet_sparse_render(pixel_coords, render_ranges, face_vertices_z, face_vertices_image, face_features,
                         knum=300, eps=1e-8):
    r"""Fully differentiable volumetric renderer devised by *Gao et al.* (DefTet, NeurIPS 2020) — reference:
    kaolin/render/mesh/deftet.py:340-417, same arguments and return values.

    Rasterizes a mesh w.r.t. a list of pixel coordinates and renders ALL the intersections within the depth
    range, sorted by depth (closest first), in padded arrays; when more than ``knum`` faces intersect a pixel
    the first ``knum`` in mesh order are kept.  Not differentiable w.r.t. ``pixel_coords``.

    Args:
        pixel_coords (torch.Tensor): :math:`(\text{batch_size}, \text{num_pixels}, 2)`.
        render_ranges (torch.Tensor): depth ranges, :math:`(\text{batch_size}, \text{num_pixels}, 2)`.
        face_vertices_z (torch.Tensor): :math:`(\text{batch_size}, \text{num_faces}, 3)`.
        face_vertices_image (torch.Tensor): :math:`(\text{batch_size}, \text{num_faces}, 3, 2)`.
        face_features (torch.Tensor or list of torch.Tensor):
            :math:`(\text{batch_size}, \text{num_faces}, 3, \text{feature_dim})` (or a list of such).
        knum (int): maximum number of faces that influence one pixel. Default: 300.
        eps (float): epsilon used to normalize barycentric weights. Default: 1e-8.

    Returns:
        (torch.Tensor or tuple of torch.Tensor, torch.LongTensor): rendered features
        :math:`(\text{batch_size}, \text{num_pixels}, \text{knum}, \text{feature_dim})` (a tuple for a list
        input) and rendered face indices (-1 is void) :math:`(\text{batch_size}, \text{num_pixels}, \text{knum})`.
    """
    is_list = isinstance(face_features, (list, tuple))
    feats = list(face_features) if is_list else [face_features]
    out = DeftetSparseRenderer.apply(pixel_coords, render_ranges, face_vertices_z, face_vertices_image, knum, eps,
                                     *feats)
    image_features, face_idx = out[:-1], out[-1]
    return (tuple(image_features) if is_list else image_features[0]), face_idx
