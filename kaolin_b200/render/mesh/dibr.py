"""Mirror of kaolin/render/mesh/dibr.py (``dibr_soft_mask``, ``dibr_rasterization``, ``DibrSoftMaskCuda``),
backed by libdibr_b200.so.

``DibrSoftMaskCuda.forward`` of the reference multiplies the coordinates, builds enlarged bboxes with five
torch kernels and saves three (B,H,W,knum) side buffers (13*knum bytes per pixel) for backward
(dibr.py:31-54).  Here one C-ABI call does setup + binning + the soft-mask kernel, writes only ``soft_mask``,
and backward re-derives each uncovered pixel's k-list with the forward arithmetic.
"""
import torch
from torch.autograd import Function

from ... import _C  # noqa: F401
from ... import _lib
from .rasterization import rasterize, _legacy_to_opengl, nvdiff, _get_nvdiff_glctx, _require_cuda  # noqa: F401

__all__ = [
    "dibr_soft_mask",
    "dibr_rasterization",
]   # = the reference's (dibr.py:21-24); dibr_rasterization_from_mesh (an addition) is exported by the package


class DibrSoftMaskCuda(Function):
    """forward(ctx, face_vertices_image, selected_face_idx, sigmainv, boxlen, knum, multiplier) -> soft_mask;
    backward(ctx, grad_soft_mask) -> (grad_face_vertices_image, None x5)   (reference: dibr.py:27-73)."""
    @staticmethod
    def forward(ctx, face_vertices_image, selected_face_idx, sigmainv, boxlen, knum, multiplier):
        fn = "dibr_soft_mask_forward_cuda"
        dev = _require_cuda(fn, face_vertices_image=face_vertices_image, selected_face_idx=selected_face_idx)
        face_vertices_image = face_vertices_image.contiguous()
        selected_face_idx = selected_face_idx.contiguous()
        if face_vertices_image.dim() != 4 or tuple(face_vertices_image.shape[2:]) != (3, 2):
            raise RuntimeError(f"{fn}: face_vertices_image must be of shape (B,F,3,2)")
        b, f = face_vertices_image.shape[0], face_vertices_image.shape[1]
        if selected_face_idx.dim() != 3 or selected_face_idx.shape[0] != b or selected_face_idx.dtype != torch.long:
            raise RuntimeError(f"{fn}: selected_face_idx must be a LongTensor of shape (B,H,W)")
        h, w = selected_face_idx.shape[1], selected_face_idx.shape[2]
        suf = _lib.suffix(face_vertices_image.dtype, fn)
        soft_mask = torch.empty((b, h, w), dtype=face_vertices_image.dtype, device=dev)
        # with a gradient to come, the forward also logs its evaluated (pixel, face) pairs behind the bins and
        # backward runs one thread per logged pair.  Records are 16 bytes and the log holds 2 per pixel (+64 Ki):
        # about 32 bytes per pixel (0.5 GB at 64 x 512^2, 2.1 GB at 64 x 1024^2) held on ctx until backward, next to
        # the bins — the reference holds 13 * knum = 390 bytes per pixel of side buffers instead.  If that allocation
        # fails the forward goes without a log (keep = 0) and backward re-derives the k-lists (recompute kernel).
        keep = 1 if ctx.needs_input_grad[0] else 0
        lib = _lib.load()
        with torch.cuda.device(dev):
            nbytes = lib.dibr_soft_mask_workspace_bytes(b, b * f, h, w, keep, face_vertices_image.element_size())
            try:
                ws = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
            except torch.cuda.OutOfMemoryError:
                if not keep:
                    raise
                keep = 0
                nbytes = lib.dibr_soft_mask_workspace_bytes(b, b * f, h, w, 0, face_vertices_image.element_size())
                ws = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
            code = getattr(lib, "dibr_soft_mask_fwd_fused_" + suf)(
                _lib.ptr(face_vertices_image), _lib.ptr(selected_face_idx), b, f, h, w, float(sigmainv),
                float(boxlen), int(knum), float(multiplier), _lib.ptr(soft_mask), 0, 0, 0, _lib.ptr(ws), nbytes,
                keep, _lib.current_stream(dev))
        _lib.check(code, fn)
        ctx.multiplier = multiplier
        ctx.sigmainv = sigmainv
        ctx.boxlen = boxlen
        ctx.knum = knum
        ctx.save_for_backward(soft_mask, face_vertices_image, selected_face_idx)
        # the workspace (bins: <= 80 bytes per face; pair log: see above) is kept for backward, which then skips the
        # setup / binning kernels; without a log nothing is kept and backward bins again
        ctx.bins = ws if keep else None
        ctx.set_materialize_grads(False)
        return soft_mask

    @staticmethod
    def backward(ctx, grad_soft_mask):
        fn = "dibr_soft_mask_backward_cuda"
        if grad_soft_mask is None:
            return None, None, None, None, None, None
        soft_mask, face_vertices_image, selected_face_idx = ctx.saved_tensors
        dev = face_vertices_image.device
        b, f = face_vertices_image.shape[0], face_vertices_image.shape[1]
        h, w = selected_face_idx.shape[1], selected_face_idx.shape[2]
        suf = _lib.suffix(face_vertices_image.dtype, fn)
        grad_soft_mask = grad_soft_mask.contiguous()
        grad_face_vertices_image = torch.empty_like(face_vertices_image)
        with torch.cuda.device(dev):
            ws, ready = ctx.bins, 2
            if ws is None:
                (ws, _), ready = _lib.workspace(b, b * f, h, w, dev), 0
            ctx.bins = None
            nbytes = ws.numel()
            code = getattr(_lib.load(), "dibr_soft_mask_bwd_fused_" + suf)(
                _lib.ptr(grad_soft_mask), _lib.ptr(soft_mask), _lib.ptr(selected_face_idx),
                _lib.ptr(face_vertices_image), b, f, h, w, float(ctx.sigmainv), float(ctx.boxlen), int(ctx.knum),
                float(ctx.multiplier), _lib.ptr(grad_face_vertices_image), _lib.ptr(ws), nbytes, ready,
                _lib.current_stream(dev))
        _lib.check(code, fn)
        return grad_face_vertices_image, None, None, None, None, None


class DibrRasterizationCuda(Function):
    """``dibr_rasterization`` with backend cuda as ONE native call each way (dibr.py:119-209 = RasterizeCuda with
    ``valid_faces = face_normals_z >= 0`` followed by DibrSoftMaskCuda): both face-list sets come from a single read of
    the coordinates, and backward accumulates the rasterizer's and the soft mask's image-plane gradients into one
    tensor.  forward(ctx, height, width, z, img, normals_z, sigmainv, boxlen, knum, multiplier, eps, *face_features)
    -> (*interpolated_features, soft_mask, face_idx); forward values are bit-identical to the two separate calls."""
    @staticmethod
    def forward(ctx, height, width, face_vertices_z, face_vertices_image, face_normals_z, sigmainv, boxlen, knum,
                multiplier, eps, *face_features):
        fn = "dibr_rasterization"
        dev = _require_cuda(fn, face_vertices_z=face_vertices_z, face_vertices_image=face_vertices_image,
                            face_normals_z=face_normals_z,
                            **{f"face_features[{i}]": f for i, f in enumerate(face_features)})
        b, f = face_vertices_z.shape[0], face_vertices_z.shape[1]
        dtype = face_vertices_z.dtype
        suf = _lib.suffix(dtype, fn)
        feats = []
        for t in face_features:
            if t.dtype != dtype or t.dim() != 4 or tuple(t.shape[:3]) != (b, f, 3):
                raise RuntimeError(f"{fn}: every face_features tensor must be (B,F,3,C) of dtype {dtype}")
            feats.append(t.contiguous())
        if face_vertices_image.dtype != dtype or tuple(face_vertices_z.shape) != (b, f, 3) or \
                tuple(face_vertices_image.shape) != (b, f, 3, 2):
            raise RuntimeError(f"{fn}: expected face_vertices_z (B,F,3), face_vertices_image (B,F,3,2)")
        if tuple(face_normals_z.shape) != (b, f) or face_normals_z.dtype != dtype:
            raise RuntimeError(f"{fn}: face_normals_z must be (B,F) of dtype {dtype}")
        img = face_vertices_image.contiguous()
        z = face_vertices_z.contiguous()
        nz = face_normals_z.contiguous()
        chans = [t.shape[-1] for t in feats]
        outs = [torch.empty((b, height, width, c), dtype=dtype, device=dev) for c in chans]
        face_idx = torch.empty((b, height, width), dtype=torch.long, device=dev)
        soft = torch.empty((b, height, width), dtype=dtype, device=dev)
        # pair log for backward: see DibrSoftMaskCuda.forward (about 32 bytes per pixel, held until backward)
        keep = 1 if ctx.needs_input_grad[3] else 0
        lib = _lib.load()
        with torch.cuda.device(dev):
            nbytes = lib.dibr_rasterization_workspace_bytes(b, b * f, int(height), int(width), keep)
            try:
                ws = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
            except torch.cuda.OutOfMemoryError:
                if not keep:
                    raise
                keep = 0
                nbytes = lib.dibr_rasterization_workspace_bytes(b, b * f, int(height), int(width), 0)
                ws = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
            code = getattr(lib, "dibr_rasterization_fwd_fused_" + suf)(
                _lib.ptr(z), _lib.ptr(img), _lib.ptr_array(feats), _lib.int_array(chans), len(feats), _lib.ptr(nz), 0,
                b, f, int(height), int(width), float(multiplier), float(eps), float(sigmainv), float(boxlen), int(knum),
                _lib.ptr_array(outs), _lib.ptr(face_idx), _lib.ptr(soft), _lib.ptr(ws), nbytes, keep,
                _lib.current_stream(dev))
        _lib.check(code, fn)
        ctx.save_for_backward(face_idx, soft, img, *feats)
        ctx.mark_non_differentiable(face_idx)
        ctx.set_materialize_grads(False)
        ctx.args = (float(multiplier), float(eps), float(sigmainv), float(boxlen), int(knum))
        ctx.out_shapes = [tuple(o.shape) for o in outs]
        ctx.bins = ws if keep else None
        return (*outs, soft, face_idx)

    @staticmethod
    def backward(ctx, *grads):
        fn = "dibr_rasterization_backward"
        g_outs, g_soft = grads[:-2], grads[-2]
        n_in = 10 + len(g_outs)
        if g_soft is None and all(g is None for g in g_outs):
            return (None,) * n_in
        face_idx, soft, img = ctx.saved_tensors[:3]
        feats = list(ctx.saved_tensors[3:])
        dev, dtype = img.device, img.dtype
        suf = _lib.suffix(dtype, fn)
        multiplier, eps, sigmainv, boxlen, knum = ctx.args
        b, f = img.shape[0], img.shape[1]
        h, w = face_idx.shape[1], face_idx.shape[2]
        any_out = any(g is not None for g in g_outs)
        gouts = [torch.zeros(shp, dtype=dtype, device=dev) if g is None else g.contiguous()
                 for g, shp in zip(g_outs, ctx.out_shapes)] if any_out else []
        g_soft = None if g_soft is None else g_soft.contiguous()
        gi = torch.empty_like(img)
        # without an upstream feature gradient the feature tensors get None (as when RasterizeCuda.backward is not run)
        gfs = [torch.empty_like(t) for t in feats] if any_out else []
        chans = [t.shape[-1] for t in feats]
        lib = _lib.load()
        with torch.cuda.device(dev):
            ws, ready = ctx.bins, 6
            ctx.bins = None
            if ws is None:
                nbytes = lib.dibr_soft_mask_workspace_bytes(b, b * f, h, w, 0, img.element_size())
                ws, ready = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev), 0
            code = getattr(lib, "dibr_rasterization_bwd_fused_" + suf)(
                _lib.ptr_array(gouts) if any_out else None, _lib.ptr(g_soft), _lib.ptr(soft), _lib.ptr(face_idx),
                _lib.ptr(img), _lib.ptr_array(feats), _lib.int_array(chans), len(feats), b, f, h, w, multiplier, eps,
                sigmainv, boxlen, knum, _lib.ptr(gi), _lib.ptr_array(gfs) if any_out else None, _lib.ptr(ws),
                ws.numel(), ready, _lib.current_stream(dev))
        _lib.check(code, fn)
        gfs = gfs if any_out else [None] * len(feats)
        return (None, None, None, gi, None, None, None, None, None, None, *gfs)


class DibrRasterizationFromMeshCuda(Function):
    """``prepare_vertices`` + ``dibr_rasterization`` as ONE native call each way (the projection
    is folded into the setup kernel, its backward runs behind the two gradient kernels inside the same call).
    forward(ctx, height, width, vertices, faces, proj, rot, trans, transform, batch, sigmainv, boxlen, knum,
    multiplier, eps, *face_features) -> (*interpolated_features, soft_mask, face_idx); gradients flow to ``vertices``,
    the camera tensors that require grad, and the face features."""
    @staticmethod
    def forward(ctx, height, width, vertices, faces, proj, rot, trans, transform, batch, sigmainv, boxlen, knum,
                multiplier, eps, *face_features):
        fn = "dibr_rasterization_from_mesh"
        dev = _require_cuda(fn, vertices=vertices, faces=faces,
                            **{f"face_features[{i}]": f for i, f in enumerate(face_features)})
        dtype = vertices.dtype
        suf = _lib.suffix(dtype, fn)
        b, f = int(batch), faces.shape[0]
        feats = []
        for t in face_features:
            if t.dtype != dtype or t.dim() != 4 or tuple(t.shape[:3]) != (b, f, 3):
                raise RuntimeError(f"{fn}: every face_features tensor must be (B,F,3,C) of dtype {dtype}")
            feats.append(t.contiguous())
        chans = [t.shape[-1] for t in feats]
        outs = [torch.empty((b, height, width, c), dtype=dtype, device=dev) for c in chans]
        face_idx = torch.empty((b, height, width), dtype=torch.long, device=dev)
        soft = torch.empty((b, height, width), dtype=dtype, device=dev)
        img = torch.empty((b, f, 3, 2), dtype=dtype, device=dev)       # written by the setup kernel
        z = torch.empty((b, f, 3), dtype=dtype, device=dev)
        need_geo = any(ctx.needs_input_grad[i] for i in (2, 4, 5, 6, 7))
        keep = 1 if need_geo else 0
        lib = _lib.load()
        with torch.cuda.device(dev):
            nbytes = lib.dibr_rasterization_workspace_bytes(b, b * f, int(height), int(width), keep)
            try:
                ws = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
            except torch.cuda.OutOfMemoryError:
                if not keep:
                    raise
                keep = 0
                nbytes = lib.dibr_rasterization_workspace_bytes(b, b * f, int(height), int(width), 0)
                ws = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
            code = getattr(lib, "dibr_rasterization_from_mesh_fwd_" + suf)(
                _lib.ptr(vertices), _lib.ptr(faces), _lib.ptr(proj), _lib.ptr(rot), _lib.ptr(trans),
                _lib.ptr(transform), vertices.shape[0], vertices.shape[1], _lib.ptr_array(feats),
                _lib.int_array(chans), len(feats), b, f, int(height), int(width), float(multiplier), float(eps),
                float(sigmainv), float(boxlen), int(knum), _lib.ptr(img), _lib.ptr(z), _lib.ptr_array(outs),
                _lib.ptr(face_idx), _lib.ptr(soft), _lib.ptr(ws), nbytes, keep, _lib.current_stream(dev))
        _lib.check(code, fn)
        ctx.save_for_backward(face_idx, soft, img, vertices, faces, proj, rot, trans, transform, *feats)
        ctx.mark_non_differentiable(face_idx)
        ctx.set_materialize_grads(False)
        ctx.args = (float(multiplier), float(eps), float(sigmainv), float(boxlen), int(knum))
        ctx.out_shapes = [tuple(o.shape) for o in outs]
        ctx.bins = ws if keep else None
        ctx.batch = b
        return (*outs, soft, face_idx)

    @staticmethod
    def backward(ctx, *grads):
        fn = "dibr_rasterization_from_mesh_backward"
        g_outs, g_soft = grads[:-2], grads[-2]
        n_in = 14 + len(g_outs)
        if g_soft is None and all(g is None for g in g_outs):
            return (None,) * n_in
        face_idx, soft, img, vertices, faces, proj, rot, trans, transform = ctx.saved_tensors[:9]
        feats = list(ctx.saved_tensors[9:])
        dev, dtype = img.device, img.dtype
        suf = _lib.suffix(dtype, fn)
        multiplier, eps, sigmainv, boxlen, knum = ctx.args
        b, f = img.shape[0], img.shape[1]
        h, w = face_idx.shape[1], face_idx.shape[2]
        need = ctx.needs_input_grad
        any_out = any(g is not None for g in g_outs)
        gouts = [torch.zeros(shp, dtype=dtype, device=dev) if g is None else g.contiguous()
                 for g, shp in zip(g_outs, ctx.out_shapes)] if any_out else []
        g_soft = None if g_soft is None else g_soft.contiguous()
        gi = torch.empty_like(img)                                    # scratch of the native call
        gfs = [torch.empty_like(t) for t in feats] if any_out else []
        chans = [t.shape[-1] for t in feats]
        g_vert = torch.empty_like(vertices) if need[2] else None
        g_proj = torch.empty_like(proj) if need[4] else None
        g_rot = torch.empty_like(rot) if rot is not None and need[5] else None
        g_trans = torch.empty_like(trans) if trans is not None and need[6] else None
        g_transform = torch.empty_like(transform) if transform is not None and need[7] else None
        geo = any(t is not None for t in (g_vert, g_proj, g_rot, g_trans, g_transform))
        lib = _lib.load()
        with torch.cuda.device(dev):
            ws, ready = ctx.bins, 6
            ctx.bins = None
            if ws is None:
                nbytes = lib.dibr_soft_mask_workspace_bytes(b, b * f, h, w, 0, img.element_size())
                ws, ready = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev), 0
            if geo:
                code = getattr(lib, "dibr_rasterization_from_mesh_bwd_" + suf)(
                    _lib.ptr_array(gouts) if any_out else None, _lib.ptr(g_soft), _lib.ptr(soft), _lib.ptr(face_idx),
                    _lib.ptr(img), _lib.ptr_array(feats), _lib.int_array(chans), len(feats), _lib.ptr(vertices),
                    _lib.ptr(faces), _lib.ptr(proj), _lib.ptr(rot), _lib.ptr(trans), _lib.ptr(transform),
                    vertices.shape[0], vertices.shape[1], b, f, h, w, multiplier, eps, sigmainv, boxlen, knum,
                    _lib.ptr(gi), _lib.ptr_array(gfs) if any_out else None, _lib.ptr(g_vert), _lib.ptr(g_proj),
                    _lib.ptr(g_rot), _lib.ptr(g_trans), _lib.ptr(g_transform), _lib.ptr(ws), ws.numel(), ready,
                    _lib.current_stream(dev))
            else:                                                     # only the face features require grad
                code = getattr(lib, "dibr_rasterization_bwd_fused_" + suf)(
                    _lib.ptr_array(gouts) if any_out else None, _lib.ptr(g_soft), _lib.ptr(soft), _lib.ptr(face_idx),
                    _lib.ptr(img), _lib.ptr_array(feats), _lib.int_array(chans), len(feats), b, f, h, w, multiplier,
                    eps, sigmainv, boxlen, knum, _lib.ptr(gi), _lib.ptr_array(gfs) if any_out else None,
                    _lib.ptr(ws), ws.numel(), ready, _lib.current_stream(dev))
        _lib.check(code, fn)
        gfs = gfs if any_out else [None] * len(feats)
        return (None, None, g_vert, None, g_proj, g_rot, g_trans, g_transform, None, None, None, None, None, None,
                *gfs)


def dibr_soft_mask(face_vertices_image, selected_face_idx,
                   sigmainv=7000, boxlen=0.02, knum=30, multiplier=1000.):
    r"""Soft foreground mask of DIB-R (reference: kaolin/render/mesh/dibr.py:75-117).

    Args:
        face_vertices_image (torch.Tensor): 2D positions of the face vertices in [-1, 1], (B, F, 3, 2).
        selected_face_idx (torch.LongTensor): rendered face index (B, H, W), 2nd output of ``rasterize``.
        sigmainv (float): smoothness term, the higher the sharper. Default: 7000.
        boxlen (float): margin over the face bounding boxes deciding which pixels a face influences. Default: 0.02.
        knum (int): maximum number of faces influencing one pixel. Default: 30.
        multiplier (float): internal enlargement of the 2D coordinates. Default: 1000.
    Returns:
        (torch.FloatTensor): the soft mask, (B, H, W).
    """
    return DibrSoftMaskCuda.apply(face_vertices_image, selected_face_idx,
                                  sigmainv, boxlen, knum, multiplier)


def dibr_rasterization(height, width, face_vertices_z, face_vertices_image,
                       face_features, face_normals_z, sigmainv=7000,
                       boxlen=0.02, knum=30, multiplier=None, eps=None,
                       rast_backend='cuda'):
    r"""DIB-R renderer: feature images, soft masks and face index maps
    (reference: kaolin/render/mesh/dibr.py:119-209; only faces with ``face_normals_z >= 0`` are drawn).

    Returns:
        (features (B,H,W,C) or tuple of tensors, soft_mask (B,H,W), face_idx (B,H,W) int64, -1 = none)
    """
    is_list = isinstance(face_features, (list, tuple))
    feats = list(face_features) if is_list else [face_features]
    if rast_backend == 'cuda' and 0 < len(feats) <= 8 and face_vertices_z.is_cuda and \
            torch.is_tensor(face_normals_z) and face_normals_z.dtype == face_vertices_z.dtype and \
            tuple(face_normals_z.shape) == tuple(face_vertices_z.shape[:2]):
        # one native call each way (same values as the composition below, bit for bit)
        *outs, soft_mask, face_idx = DibrRasterizationCuda.apply(
            height, width, face_vertices_z, face_vertices_image, face_normals_z, sigmainv, boxlen, knum,
            1000 if multiplier is None else multiplier, 1e-8 if eps is None else eps, *feats)
        return (tuple(outs) if is_list else outs[0]), soft_mask, face_idx
    interpolated_features, face_idx = rasterize(
        height, width,
        face_vertices_z,
        face_vertices_image,
        face_features,
        face_normals_z >= 0.,
        multiplier,
        eps,
        rast_backend
    )
    _multiplier = 1000. if multiplier is None else multiplier
    soft_mask = dibr_soft_mask(
        face_vertices_image,
        face_idx,
        sigmainv,
        boxlen,
        knum,
        _multiplier
    )
    return interpolated_features, soft_mask, face_idx


def dibr_rasterization_from_mesh(height, width, vertices, faces, face_features, camera_proj, camera_rot=None,
                                 camera_trans=None, camera_transform=None, sigmainv=7000, boxlen=0.02, knum=30,
                                 multiplier=None, eps=None):
    r"""``prepare_vertices`` followed by ``dibr_rasterization`` in one native call each way — the composition

    .. code-block:: python

        cam, img, nrm = prepare_vertices(vertices, faces, camera_proj, camera_rot, camera_trans, camera_transform)
        dibr_rasterization(height, width, cam[..., 2], img, face_features, nrm[..., 2], sigmainv, boxlen, knum,
                           multiplier, eps)

    of the reference (kaolin/render/mesh/utils.py:128-175, dibr.py:119-209; the training step of
    examples/tutorial/ian_dibr.py:246-252), with identical results: the setup kernel projects the faces itself, the
    camera-space (B,F,3,3) tensor and the normals are never written, and backward returns ``vertices.grad`` (and the
    camera gradients) directly.  Arguments as in the two reference functions; ``face_features`` is (B,F,3,C) or a list.

    Returns:
        (features (B,H,W,C) or tuple of tensors, soft_mask (B,H,W), face_idx (B,H,W) int64, -1 = none)
    """
    from .utils import _mesh_and_cameras
    is_list = isinstance(face_features, (list, tuple))
    feats = list(face_features) if is_list else [face_features]
    if not 0 < len(feats) <= 8:
        raise ValueError("dibr_rasterization_from_mesh: between 1 and 8 face feature tensors are supported")
    v, f, proj, rot, trans, transform, batch = _mesh_and_cameras(vertices, faces, camera_proj, camera_rot,
                                                                 camera_trans, camera_transform)
    *outs, soft_mask, face_idx = DibrRasterizationFromMeshCuda.apply(
        height, width, v, f, proj, rot, trans, transform, batch, sigmainv, boxlen, knum,
        1000 if multiplier is None else multiplier, 1e-8 if eps is None else eps, *feats)
    return (tuple(outs) if is_list else outs[0]), soft_mask, face_idx
