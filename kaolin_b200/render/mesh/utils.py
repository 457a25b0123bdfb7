"""Mirror of ``kaolin.render.mesh.utils.prepare_vertices`` (kaolin/render/mesh/utils.py:128-175), the producer
of the rasterization path's inputs, and of ``texture_mapping`` (utils.py:23-76), the consumer of its
interpolated uv map, backed by libdibr_b200.so.

The reference composes rotate_translate_points / ``padded @ camera_transform``, perspective_camera, two
index_vertices_by_faces gathers and face_normals(unit=True) out of ~12 torch kernels and materialises the
per-vertex camera / image coordinates and two expanded gathers on the way.  Here ONE kernel per direction:
thread = (camera, face) gathers its three vertices, transforms, projects, and writes the three outputs;
backward scatters into ``vertices.grad`` (summed over cameras when one mesh is broadcast over the batch).

Every floating-point input is differentiable, as in the reference (pure torch): ``vertices`` and, when they
require grad, ``camera_proj`` / ``camera_rot`` / ``camera_trans`` / ``camera_transform`` (warp-reduced sums in
the same backward kernel).
"""
import torch
from torch.autograd import Function

from ... import _lib

__all__ = ["prepare_vertices", "texture_mapping"]


def _cam_ptrs(rot, trans, transform):
    return _lib.ptr(rot), _lib.ptr(trans), _lib.ptr(transform)


class PrepareVerticesCuda(Function):
    """forward(ctx, vertices (Bv,V,3), faces (F,3), proj (3), rot|None, trans|None, transform|None, B)
    -> face_vertices_camera (B,F,3,3), face_vertices_image (B,F,3,2), face_normals (B,F,3)."""
    @staticmethod
    def forward(ctx, vertices, faces, proj, rot, trans, transform, batch):
        fn = "prepare_vertices"
        dev = vertices.device
        suf = _lib.suffix(vertices.dtype, fn)
        bv, v = vertices.shape[0], vertices.shape[1]
        f = faces.shape[0]
        fcam = torch.empty((batch, f, 3, 3), dtype=vertices.dtype, device=dev)
        fimg = torch.empty((batch, f, 3, 2), dtype=vertices.dtype, device=dev)
        fnrm = torch.empty((batch, f, 3), dtype=vertices.dtype, device=dev)
        with torch.cuda.device(dev):
            code = getattr(_lib.load(), "dibr_prepare_vertices_fwd_" + suf)(
                _lib.ptr(vertices), _lib.ptr(faces), _lib.ptr(proj), *_cam_ptrs(rot, trans, transform),
                batch, bv, v, f, _lib.ptr(fcam), _lib.ptr(fimg), _lib.ptr(fnrm), _lib.current_stream(dev))
        _lib.check(code, fn)
        ctx.save_for_backward(vertices, faces, proj, rot, trans, transform)
        ctx.batch = batch
        ctx.set_materialize_grads(False)
        return fcam, fimg, fnrm

    @staticmethod
    def backward(ctx, g_cam, g_img, g_nrm):
        fn = "prepare_vertices_backward"
        vertices, faces, proj, rot, trans, transform = ctx.saved_tensors
        need = ctx.needs_input_grad
        if not any(need[:1] + need[2:6]):
            return (None,) * 7
        dev = vertices.device
        suf = _lib.suffix(vertices.dtype, fn)
        g_cam = None if g_cam is None else g_cam.contiguous()
        g_img = None if g_img is None else g_img.contiguous()
        g_nrm = None if g_nrm is None else g_nrm.contiguous()
        grad_vertices = torch.empty_like(vertices) if need[0] else None
        g_proj = torch.empty_like(proj) if need[2] else None
        g_rot = torch.empty_like(rot) if rot is not None and need[3] else None
        g_trans = torch.empty_like(trans) if trans is not None and need[4] else None
        g_transform = torch.empty_like(transform) if transform is not None and need[5] else None
        with torch.cuda.device(dev):
            code = getattr(_lib.load(), "dibr_prepare_vertices_bwd_" + suf)(
                _lib.ptr(g_cam), _lib.ptr(g_img), _lib.ptr(g_nrm), _lib.ptr(vertices), _lib.ptr(faces),
                _lib.ptr(proj), *_cam_ptrs(rot, trans, transform), ctx.batch, vertices.shape[0],
                vertices.shape[1], faces.shape[0], _lib.ptr(grad_vertices), _lib.ptr(g_proj), _lib.ptr(g_rot),
                _lib.ptr(g_trans), _lib.ptr(g_transform), _lib.current_stream(dev))
        _lib.check(code, fn)
        return grad_vertices, None, g_proj, g_rot, g_trans, g_transform, None


def _camera_tensor(name, t, shape_tail, like):
    if t.dim() != len(shape_tail) + 1 or tuple(t.shape[1:]) != shape_tail:
        raise RuntimeError(f"prepare_vertices: {name} must be of shape (batch_size, "
                           f"{', '.join(str(s) for s in shape_tail)}), got {tuple(t.shape)}")
    if t.device != like.device:
        raise RuntimeError(f"prepare_vertices: {name} must be on {like.device}, got {t.device}")
    if t.dtype != like.dtype:
        raise RuntimeError(f"prepare_vertices: {name} must be of dtype {like.dtype}, got {t.dtype}")
    return t.contiguous()


def prepare_vertices(vertices, faces, camera_proj, camera_rot=None, camera_trans=None, camera_transform=None):
    r"""Move and project vertices to cameras, then index them with faces (reference:
    kaolin/render/mesh/utils.py:128-175 — same arguments, same three return values).

    Args:
        vertices (torch.Tensor): of shape :math:`(\text{batch_size}, \text{num_vertices}, 3)`; a batch of 1 is
            broadcast over the cameras' batch, as torch broadcasting does in the reference.
        faces (torch.LongTensor): of shape :math:`(\text{num_faces}, 3)`.
        camera_proj (torch.Tensor): projection vector, of shape :math:`(3, 1)`.
        camera_rot (torch.Tensor, optional): of shape :math:`(\text{batch_size}, 3, 3)`.
        camera_trans (torch.Tensor, optional): of shape :math:`(\text{batch_size}, 3)`.
        camera_transform (torch.Tensor, optional): of shape :math:`(\text{batch_size}, 4, 3)`; replaces
            ``camera_rot`` and ``camera_trans``.

    Returns:
        (torch.Tensor, torch.Tensor, torch.Tensor): face_vertices_camera :math:`(B, F, 3, 3)`,
        face_vertices_image :math:`(B, F, 3, 2)`, face_normals :math:`(B, F, 3)`.
    """
    return PrepareVerticesCuda.apply(*_mesh_and_cameras(vertices, faces, camera_proj, camera_rot, camera_trans,
                                                        camera_transform))


def _mesh_and_cameras(vertices, faces, camera_proj, camera_rot=None, camera_trans=None, camera_transform=None):
    """Argument checks and broadcasting of prepare_vertices (utils.py:128-175); returns the contiguous tensors
    (vertices, faces, proj (3), rot | None, trans | None, transform | None, batch) the native calls take."""
    if camera_transform is None:
        assert camera_trans is not None and camera_rot is not None, \
            "camera_transform or camera_trans and camera_rot must be defined"
    else:
        assert camera_trans is None and camera_rot is None, \
            "camera_trans and camera_rot must be None when camera_transform is defined"
    if not vertices.is_cuda:
        raise RuntimeError("prepare_vertices: vertices must be a CUDA tensor (libdibr_b200.so has no CPU path)")
    if vertices.dim() != 3 or vertices.shape[-1] != 3:
        raise RuntimeError(f"prepare_vertices: vertices must be of shape (batch_size, num_vertices, 3), "
                           f"got {tuple(vertices.shape)}")
    if faces.dim() != 2 or faces.dtype != torch.long:
        raise RuntimeError("prepare_vertices: faces must be a LongTensor of shape (num_faces, face_size)")
    if faces.shape[1] != 3:
        # kaolin/ops/mesh/trianglemesh.py:326-327
        raise NotImplementedError("face_normals is only implemented for triangle meshes")
    if faces.device != vertices.device:
        raise RuntimeError(f"prepare_vertices: faces must be on {vertices.device}, got {faces.device}")
    if camera_proj.numel() != 3:
        raise RuntimeError(f"prepare_vertices: camera_proj must hold 3 values, got {tuple(camera_proj.shape)}")
    proj = camera_proj.to(device=vertices.device, dtype=vertices.dtype).reshape(3).contiguous()
    rot = trans = transform = None
    if camera_transform is None:
        rot = _camera_tensor("camera_rot", camera_rot, (3, 3), vertices)
        # legacy.py:35 views camera_trans as (-1, 1, 3)
        trans = _camera_tensor("camera_trans", camera_trans.reshape(-1, 3), (3,), vertices)
        batch = rot.shape[0]
        if trans.shape[0] != batch:
            if trans.shape[0] != 1:
                raise RuntimeError("prepare_vertices: camera_rot and camera_trans batch sizes differ")
            trans = trans.expand(batch, 3).contiguous()
    else:
        transform = _camera_tensor("camera_transform", camera_transform, (4, 3), vertices)
        batch = transform.shape[0]
    bv = vertices.shape[0]
    if batch == 1 and bv > 1:                         # one camera broadcast over a batch of meshes
        batch = bv
        if transform is None:
            rot, trans = rot.expand(batch, 3, 3).contiguous(), trans.expand(batch, 3).contiguous()
        else:
            transform = transform.expand(batch, 4, 3).contiguous()
    if bv != 1 and bv != batch:
        raise RuntimeError(f"prepare_vertices: vertices batch ({bv}) and camera batch ({batch}) do not broadcast")
    return vertices.contiguous(), faces.contiguous(), proj, rot, trans, transform, batch


# ----------------------------------------------------------------------------------------------------------
_TEX_MODES = {"nearest": 0, "bilinear": 1}


class TextureMappingCuda(Function):
    """forward(ctx, uv (B,N,2), texture_maps (Bt,C,Ht,Wt), mask (B,N)|None, mode) -> (B,N,C)."""
    @staticmethod
    def forward(ctx, uv, texture_maps, mask, mode):
        fn = "texture_mapping"
        dev = uv.device
        suf = _lib.suffix(uv.dtype, fn)
        b, n = uv.shape[0], uv.shape[1]
        bt, c, ht, wt = texture_maps.shape
        out = torch.empty((b, n, c), dtype=uv.dtype, device=dev)
        with torch.cuda.device(dev):
            code = getattr(_lib.load(), "dibr_texture_mapping_fwd_" + suf)(
                _lib.ptr(uv), _lib.ptr(texture_maps), _lib.ptr(mask), b, n, bt, c, ht, wt, mode, _lib.ptr(out),
                _lib.current_stream(dev))
        _lib.check(code, fn)
        ctx.save_for_backward(uv, texture_maps, mask)
        ctx.mode = mode
        return out

    @staticmethod
    def backward(ctx, grad_out):
        fn = "texture_mapping_backward"
        uv, texture_maps, mask = ctx.saved_tensors
        dev = uv.device
        suf = _lib.suffix(uv.dtype, fn)
        need_uv, need_tex, need_mask = ctx.needs_input_grad[0], ctx.needs_input_grad[1], \
            mask is not None and ctx.needs_input_grad[2]
        g_uv = torch.empty_like(uv) if need_uv else None
        g_tex = torch.empty_like(texture_maps) if need_tex else None
        g_mask = torch.empty_like(mask) if need_mask else None
        b, n = uv.shape[0], uv.shape[1]
        bt, c, ht, wt = texture_maps.shape
        with torch.cuda.device(dev):
            code = getattr(_lib.load(), "dibr_texture_mapping_bwd_" + suf)(
                _lib.ptr(grad_out.contiguous()), _lib.ptr(uv), _lib.ptr(texture_maps), _lib.ptr(mask), b, n, bt, c,
                ht, wt, ctx.mode, _lib.ptr(g_uv), _lib.ptr(g_tex), _lib.ptr(g_mask), _lib.current_stream(dev))
        _lib.check(code, fn)
        return g_uv, g_tex, g_mask, None


def texture_mapping(texture_coordinates, texture_maps, mode='nearest', mask=None):
    r"""Interpolates texture_maps by dense or sparse texture_coordinates (reference:
    kaolin/render/mesh/utils.py:23-76 — same arguments and return value).

    Args:
        texture_coordinates (torch.FloatTensor): dense :math:`(\text{batch_size}, h, w, 2)` or sparse
            :math:`(\text{batch_size}, \text{num_points}, 2)`, OpenGL convention in [0, 1] (values outside
            are clamped, y goes up).
        texture_maps (torch.FloatTensor): :math:`(\text{batch_size}, \text{num_channels}, h', w')`.
            Extension: a batch of 1 is shared by all ``batch_size`` coordinate maps (no ``repeat`` needed).
        mode (str): ``'nearest'`` or ``'bilinear'`` (``torch.nn.functional.grid_sample`` with
            ``align_corners=False, padding_mode='border'``).
        mask (torch.FloatTensor, optional): extension — :math:`(\text{batch_size}, h, w[, 1])`; when given the
            result is ``torch.clamp(texture * mask, 0., 1.)`` computed in the same kernel (the epilogue of
            examples/tutorial/ian_dibr.py:254-258).

    Returns:
        (torch.FloatTensor): :math:`(\text{batch_size}, h, w, \text{num_channels})` or
        :math:`(\text{batch_size}, \text{num_points}, \text{num_channels})`.
    """
    if mode not in _TEX_MODES:
        # torch.nn.functional.grid_sample raises ValueError for an unknown mode
        raise ValueError(f"nn.functional.grid_sample(): expected mode to be 'bilinear', 'nearest' or "
                         f"'bicubic', but got: '{mode}'" if mode != 'bicubic' else
                         "texture_mapping: mode 'bicubic' is not implemented by libdibr_b200.so")
    if not texture_coordinates.is_cuda:
        raise RuntimeError("texture_mapping: texture_coordinates must be a CUDA tensor "
                           "(libdibr_b200.so has no CPU path)")
    if texture_coordinates.dim() not in (3, 4) or texture_coordinates.shape[-1] != 2:
        raise RuntimeError(f"texture_mapping: texture_coordinates must be of shape (batch_size, h, w, 2) or "
                           f"(batch_size, num_points, 2), got {tuple(texture_coordinates.shape)}")
    if texture_maps.dim() != 4:
        raise RuntimeError(f"texture_mapping: texture_maps must be of shape (batch_size, num_channels, h', w'), "
                           f"got {tuple(texture_maps.shape)}")
    if texture_maps.device != texture_coordinates.device or texture_maps.dtype != texture_coordinates.dtype:
        raise RuntimeError("texture_mapping: texture_maps and texture_coordinates must share device and dtype")
    batch_size = texture_coordinates.shape[0]
    if texture_maps.shape[0] not in (1, batch_size):
        raise RuntimeError(f"texture_mapping: texture_maps batch ({texture_maps.shape[0]}) does not match "
                           f"texture_coordinates batch ({batch_size})")
    num_channels = texture_maps.shape[1]
    uv = texture_coordinates.reshape(batch_size, -1, 2).contiguous()
    m = None
    if mask is not None:
        if mask.numel() != uv.shape[0] * uv.shape[1]:
            raise RuntimeError(f"texture_mapping: mask of shape {tuple(mask.shape)} does not match the "
                               f"texture_coordinates {tuple(texture_coordinates.shape)}")
        m = mask.to(uv.dtype).reshape(batch_size, -1).contiguous()
    out = TextureMappingCuda.apply(uv, texture_maps.contiguous(), m, _TEX_MODES[mode])
    return out.reshape(batch_size, *texture_coordinates.shape[1:-1], num_channels)
