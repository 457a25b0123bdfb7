"""Mirror of kaolin/render/mesh/rasterization.py (public names, signatures, defaults, warnings, errors and
autograd.Function forward/backward signatures), backed by libdibr_b200.so.

What changed underneath: ``RasterizeCuda.forward`` no longer runs ~15 torch kernels (valid-face packing,
x multiplier, bboxes, index remapping — rasterization.py:309-363 of the reference) around the native op; a
single C-ABI call does setup, tile binning, the depth-tested face search and the interpolation, and writes
``face_idx`` already as original face indices.  The barycentric weights are not saved: backward
recomputes them bit-exactly.
"""
import logging
import warnings

import torch
from torch.autograd import Function

from ... import _C  # noqa: F401  (kept importable like `from kaolin import _C`)
from ... import _lib

# reference: rasterization.py:25-33 tries `import nvdiffrast.torch`.  The OpenGL backends are unavailable on this
# box (no graphics stack); the names stay importable and every nvdiffrast entry raises the reference's ValueError.
_has_nvdiffrast = False
nvdiff = None
logging.getLogger(__name__).info("Cannot import nvdiffrast")

__all__ = [
    'rasterize',
]

_NVDIFF_MSG = ("nvdiffrast must be installed to be used as backend, but failed to import. "
               "See https://nvlabs.github.io/nvdiffrast/#installation for installation instructions.")


def _get_nvdiff_glctx(device):
    """reference: rasterization.py:35-39.  OpenGL contexts are unavailable on this box (no graphics stack):
    the name stays importable (dibr.py:20 of the reference imports it) and raises the reference's error."""
    raise ValueError(_NVDIFF_MSG)


def _legacy_to_opengl(face_vertices_image, face_vertices_z, valid_faces=None):
    """reference: rasterization.py:46-82 (feeds the nvdiffrast backends only) — unavailable on this box."""
    raise ValueError(_NVDIFF_MSG)


def _require_cuda(fn, **tensors):
    dev = None
    for name, t in tensors.items():
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError(f"Expected tensor for argument '{name}' to have CUDA type; but got {t.device.type} "
                               f"instead (while checking arguments for {fn})")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"Expected tensor for argument '{name}' to be on GPU {dev.index} (while checking "
                               f"arguments for {fn})")
    return dev


def _rasterize_backward(grad_interpolated_features, face_idx, output_weights, face_vertices_image, face_features,
                        multiplier, eps):
    fn = "rasterize_backward_cuda"
    dev = face_vertices_image.device
    b, h, w, c = grad_interpolated_features.shape
    f = face_vertices_image.shape[1]
    suf = _lib.suffix(face_vertices_image.dtype, fn)
    gi = torch.empty_like(face_vertices_image)
    gf = torch.empty_like(face_features)
    with torch.cuda.device(dev):
        code = getattr(_lib.load(), "dibr_rasterize_backward_" + suf)(
            _lib.ptr(grad_interpolated_features), _lib.ptr(face_idx), _lib.ptr(output_weights),
            _lib.ptr(face_vertices_image), _lib.ptr(face_features), b, f, c, h, w, float(multiplier), float(eps),
            _lib.ptr(gi), _lib.ptr(gf), _lib.current_stream(dev))
    _lib.check(code, fn)
    return gi, gf


class NvdiffRasterizeFwdCudaBwd(Function):
    """backend 'nvdiffrast_fwd' (reference: rasterization.py:145-241): OpenGL forward through nvdiffrast, this
    library's backward.  The forward needs OpenGL — unavailable on this box — and raises the reference's
    ValueError; the backward is the same native op as RasterizeCuda's (rasterization.py:230-238)."""
    @staticmethod
    def forward(ctx, height, width, face_vertices_z, face_vertices_image, face_features, valid_faces, eps):
        raise ValueError(_NVDIFF_MSG)

    @staticmethod
    def backward(ctx, grad_interpolated_features, grad_face_idx):
        face_idx, output_weights, face_vertices_image, face_features = ctx.saved_tensors
        gi, gf = _rasterize_backward(grad_interpolated_features.contiguous(), face_idx, output_weights,
                                     face_vertices_image, face_features, 1.0, ctx.eps)
        return None, None, None, gi, gf, None, None


def _nvdiff_rasterize(height, width, face_vertices_z, face_vertices_image, face_features, valid_faces):
    # reference: rasterization.py:84-143 — needs OpenGL: unavailable on this box
    raise ValueError(_NVDIFF_MSG)


class RasterizeCuda(Function):
    """torch.autograd.Function for ``rasterize`` with backend cuda (reference: rasterization.py:243-388).

    forward(ctx, height, width, face_vertices_z, face_vertices_image, face_features, valid_faces,
    multiplier, eps) -> (interpolated_features, face_idx); backward returns gradients for
    face_vertices_image and face_features only (no gradient flows to face_vertices_z)."""
    @staticmethod
    def forward(ctx, height, width, face_vertices_z, face_vertices_image, face_features, valid_faces,
                multiplier, eps):
        fn = "packed_rasterize_forward_cuda"
        dev = _require_cuda(fn, face_vertices_z=face_vertices_z, face_vertices_image=face_vertices_image,
                            face_features=face_features, valid_faces=valid_faces)
        batch_size, num_faces = face_vertices_z.shape[0], face_vertices_z.shape[1]
        feat_dim = face_features.shape[-1]
        dtype = face_vertices_z.dtype
        suf = _lib.suffix(dtype, fn)
        if face_vertices_image.dtype != dtype or face_features.dtype != dtype:
            raise RuntimeError(f"{fn}: all floating-point inputs must have dtype {dtype}")
        if tuple(face_vertices_z.shape) != (batch_size, num_faces, 3) or \
                tuple(face_vertices_image.shape) != (batch_size, num_faces, 3, 2) or \
                tuple(face_features.shape) != (batch_size, num_faces, 3, feat_dim):
            raise RuntimeError(f"{fn}: expected face_vertices_z (B,F,3), face_vertices_image (B,F,3,2), "
                               "face_features (B,F,3,C)")
        face_features = face_features.contiguous()
        face_vertices_image = face_vertices_image.contiguous()
        face_vertices_z = face_vertices_z.contiguous()
        valid_u8 = None
        if valid_faces is not None:
            if tuple(valid_faces.shape) != (batch_size, num_faces):
                raise RuntimeError(f"{fn}: valid_faces must be of shape (B,F)")
            valid_u8 = valid_faces.to(torch.bool).contiguous().view(torch.uint8)
        interpolated_features = torch.empty((batch_size, height, width, feat_dim), dtype=dtype, device=dev)
        face_idx = torch.empty((batch_size, height, width), dtype=torch.long, device=dev)
        with torch.cuda.device(dev):
            ws, nbytes = _lib.workspace(batch_size, batch_size * num_faces, height, width, dev)
            code = getattr(_lib.load(), "dibr_rasterize_fwd_fused_" + suf)(
                _lib.ptr(face_vertices_z), _lib.ptr(face_vertices_image), _lib.ptr(face_features),
                _lib.ptr(valid_u8), batch_size, num_faces, feat_dim, int(height), int(width), float(multiplier),
                float(eps), _lib.ptr(interpolated_features), _lib.ptr(face_idx), 0, _lib.ptr(ws), nbytes,
                _lib.current_stream(dev))
        _lib.check(code, fn)
        ctx.save_for_backward(face_idx, face_vertices_image, face_features)
        ctx.mark_non_differentiable(face_idx)
        ctx.set_materialize_grads(False)      # an unused output costs nothing (no zero-filled (B,H,W,C) gradient)
        ctx.eps = eps
        ctx.multiplier = multiplier
        return interpolated_features, face_idx

    @staticmethod
    def backward(ctx, grad_interpolated_features, grad_face_idx):
        if grad_interpolated_features is None:
            return None, None, None, None, None, None, None, None, None
        face_idx, face_vertices_image, face_features = ctx.saved_tensors
        gi, gf = _rasterize_backward(grad_interpolated_features.contiguous(), face_idx, None,
                                     face_vertices_image, face_features, ctx.multiplier, ctx.eps)
        return None, None, None, gi, gf, None, None, None, None


class RasterizeCudaList(Function):
    """``rasterize`` with backend cuda for a LIST of feature tensors, without the reference's
    ``torch.cat`` (rasterization.py:480-481) nor its output slicing (:499-505): the kernels read the n
    tensors in place and write n separate outputs, and backward receives n gradients directly.
    forward(ctx, height, width, z, img, valid_faces, multiplier, eps, *face_features) ->
    (*interpolated_features, face_idx).  Values are identical to RasterizeCuda on the concatenation."""
    @staticmethod
    def forward(ctx, height, width, face_vertices_z, face_vertices_image, valid_faces, multiplier, eps,
                *face_features):
        fn = "packed_rasterize_forward_cuda"
        dev = _require_cuda(fn, face_vertices_z=face_vertices_z, face_vertices_image=face_vertices_image,
                            valid_faces=valid_faces, **{f"face_features[{i}]": f for i, f in enumerate(face_features)})
        batch_size, num_faces = face_vertices_z.shape[0], face_vertices_z.shape[1]
        dtype = face_vertices_z.dtype
        suf = _lib.suffix(dtype, fn)
        if len(face_features) > 8:
            raise RuntimeError(f"{fn}: at most 8 feature tensors per call")
        feats = []
        for f in face_features:
            if f.dtype != dtype or f.dim() != 4 or tuple(f.shape[:3]) != (batch_size, num_faces, 3):
                raise RuntimeError(f"{fn}: every face_features tensor must be (B,F,3,C) of dtype {dtype}")
            feats.append(f.contiguous())
        if face_vertices_image.dtype != dtype or tuple(face_vertices_z.shape) != (batch_size, num_faces, 3) or \
                tuple(face_vertices_image.shape) != (batch_size, num_faces, 3, 2):
            raise RuntimeError(f"{fn}: expected face_vertices_z (B,F,3), face_vertices_image (B,F,3,2)")
        face_vertices_image = face_vertices_image.contiguous()
        face_vertices_z = face_vertices_z.contiguous()
        valid_u8 = None
        if valid_faces is not None:
            if tuple(valid_faces.shape) != (batch_size, num_faces):
                raise RuntimeError(f"{fn}: valid_faces must be of shape (B,F)")
            valid_u8 = valid_faces.to(torch.bool).contiguous().view(torch.uint8)
        chans = [f.shape[-1] for f in feats]
        outs = [torch.empty((batch_size, height, width, c), dtype=dtype, device=dev) for c in chans]
        face_idx = torch.empty((batch_size, height, width), dtype=torch.long, device=dev)
        with torch.cuda.device(dev):
            ws, nbytes = _lib.workspace(batch_size, batch_size * num_faces, height, width, dev)
            code = getattr(_lib.load(), "dibr_rasterize_fwd_multi_" + suf)(
                _lib.ptr(face_vertices_z), _lib.ptr(face_vertices_image), _lib.ptr_array(feats),
                _lib.int_array(chans), len(feats), _lib.ptr(valid_u8), batch_size, num_faces, int(height),
                int(width), float(multiplier), float(eps), _lib.ptr_array(outs), _lib.ptr(face_idx), 0,
                _lib.ptr(ws), nbytes, _lib.current_stream(dev))
        _lib.check(code, fn)
        ctx.save_for_backward(face_idx, face_vertices_image, *feats)
        ctx.mark_non_differentiable(face_idx)
        ctx.set_materialize_grads(False)      # gradients of unused outputs arrive as None
        ctx.eps = eps
        ctx.multiplier = multiplier
        ctx.out_shapes = [tuple(o.shape) for o in outs]
        return (*outs, face_idx)

    @staticmethod
    def backward(ctx, *grads):
        fn = "rasterize_backward_cuda"
        if all(g is None for g in grads[:-1]):
            return (None,) * (7 + len(grads) - 1)
        face_idx, face_vertices_image = ctx.saved_tensors[:2]
        feats = list(ctx.saved_tensors[2:])
        dev = face_vertices_image.device
        dtype = face_vertices_image.dtype
        suf = _lib.suffix(dtype, fn)
        gouts = [torch.zeros(shp, dtype=dtype, device=dev) if g is None else g.contiguous()
                 for g, shp in zip(grads[:-1], ctx.out_shapes)]
        b, f = face_vertices_image.shape[0], face_vertices_image.shape[1]
        h, w = face_idx.shape[1], face_idx.shape[2]
        gi = torch.empty_like(face_vertices_image)
        gfs = [torch.empty_like(t) for t in feats]
        chans = [t.shape[-1] for t in feats]
        with torch.cuda.device(dev):
            code = getattr(_lib.load(), "dibr_rasterize_backward_multi_" + suf)(
                _lib.ptr_array(gouts), _lib.ptr(face_idx), 0, _lib.ptr(face_vertices_image), _lib.ptr_array(feats),
                _lib.int_array(chans), len(feats), b, f, h, w, float(ctx.multiplier), float(ctx.eps), _lib.ptr(gi),
                _lib.ptr_array(gfs), _lib.current_stream(dev))
        _lib.check(code, fn)
        return (None, None, None, gi, None, None, None, *gfs)


def rasterize(height,
              width,
              face_vertices_z,
              face_vertices_image,
              face_features,
              valid_faces=None,
              multiplier=None,
              eps=None,
              backend='cuda'):
    r"""Fully differentiable rasterization: renders triangle meshes with per-vertex per-face features to
    feature images (reference: kaolin/render/mesh/rasterization.py:390-506, same arguments and returns).

    Args:
        height, width (int): size of the rendered images.
        face_vertices_z (torch.FloatTensor): depth of the face vertices in camera space, (B, F, 3).
        face_vertices_image (torch.FloatTensor): 2D positions on the image plane in [-1, 1], (B, F, 3, 2).
        face_features (torch.FloatTensor or list of torch.FloatTensor): (B, F, 3, C) or a list of such.
        valid_faces (torch.BoolTensor): mask of faces being rasterized, (B, F). Default: all.
        multiplier (int): coordinate enlargement used by the 'cuda' backend. Default: 1000.
        eps (float): epsilon used to normalise barycentric weights. Default: 1e-8.
        backend (string): one of ['cuda', 'nvdiffrast', 'nvdiffrast_fwd'].

    Returns:
        (torch.FloatTensor or tuple, torch.LongTensor): features (B, H, W, C) — a tuple of tensors if
        ``face_features`` was a list — and the face index map (B, H, W), -1 where no face is rendered.
    """
    if multiplier is None:
        multiplier = 1000
    elif backend in ['nvdiffrast', 'nvdiffrast_fwd']:
        warnings.warn(f'in "rasterize": multiplier is ignored with backend "{backend}"',
                      UserWarning)
    if eps is None:
        eps = 1e-8
    elif backend == 'nvdiffrast':
        warnings.warn(f'in "rasterize": eps is ignored with backend "{backend}"',
                      UserWarning)
    batch_size, num_faces, _ = face_vertices_z.shape
    if backend == 'cuda' and isinstance(face_features, (list, tuple)) and 0 < len(face_features) <= 8:
        # list of features: consumed in place, no concatenation / slicing (see RasterizeCudaList)
        *image_features, face_idx = RasterizeCudaList.apply(
            height, width, face_vertices_z, face_vertices_image, valid_faces, multiplier, eps, *face_features)
        return tuple(image_features), face_idx
    _face_features = torch.cat(face_features, dim=-1) \
        if isinstance(face_features, (list, tuple)) else face_features

    if backend == 'nvdiffrast':
        image_features, face_idx = _nvdiff_rasterize(
            height, width, face_vertices_z, face_vertices_image,
            _face_features, valid_faces)
    elif backend == 'nvdiffrast_fwd':
        image_features, face_idx = NvdiffRasterizeFwdCudaBwd.apply(
            height, width, face_vertices_z, face_vertices_image,
            _face_features, valid_faces, eps)
    elif backend == 'cuda':
        image_features, face_idx = RasterizeCuda.apply(
            height, width, face_vertices_z, face_vertices_image,
            _face_features, valid_faces, multiplier, eps)
    else:
        raise ValueError(f'"{backend}" is not a valid backend, ',
                         'valid choices are ["cuda", "nvdiffrast", "nvdiffrast_fwd"]')

    if isinstance(face_features, (list, tuple)):
        _image_features = []
        cur_idx = 0
        for face_feature in face_features:
            _image_features.append(image_features[..., cur_idx:cur_idx + face_feature.shape[-1]])
            cur_idx += face_feature.shape[-1]
        image_features = tuple(_image_features)
    return image_features, face_idx
