"""ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path (kaolin_b200 never imports this).

ctypes front-end of oracle/dibr_oracle.cpp plus a CPU restatement of the torch glue that surrounds
the four reference ops:

* ``OracleC.render.mesh.*``  — same four functions, argument order and return conventions as the
  reference's ``kaolin._C.render.mesh`` (kaolin/csrc/bindings.cpp:75-80), on CPU tensors.
* ``rasterize`` / ``dibr_soft_mask`` / ``dibr_rasterization`` — restate
  kaolin/render/mesh/rasterization.py:290-369,390-506 and kaolin/render/mesh/dibr.py:27-209
  (forward only, plus explicit backward helpers) on top of ``OracleC``.

Parity status: pinned (see header of dibr_oracle.cpp).
"""
import ctypes
import os
import types

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "_ref", "libdibr_oracle.so")
        if not os.path.exists(path):
            from . import build as _b
            _b.build_oracle()
        _LIB = ctypes.CDLL(path)
        _LIB.oracle_num_threads.restype = ctypes.c_int
        _LIB.oracle_check_eps_add_equivalence.restype = ctypes.c_int64
    return _LIB


def _suf(t):
    if t.dtype == torch.float32:
        return "f32"
    if t.dtype == torch.float64:
        return "f64"
    raise RuntimeError(f"oracle not implemented for '{t.dtype}'")


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _chk(*ts):
    for t in ts:
        assert t.device.type == "cpu" and t.is_contiguous()


def num_threads():
    return lib().oracle_num_threads()


def set_num_threads(n):
    lib().oracle_set_num_threads(ctypes.c_int(int(n)))


# ------------------------------------------------------------------ _C-shaped functions (CPU tensors)
def packed_rasterize_forward_cuda(height, width, face_vertices_z, face_vertices_image, face_bboxes,
                                  face_features, first_idx_face_per_mesh, multiplier, eps,
                                  rows=None):
    """rasterization.cpp:49-104 (checks + allocation) and rasterization_cuda.cu:43-192."""
    _chk(face_vertices_z, face_vertices_image, face_bboxes, face_features, first_idx_face_per_mesh)
    n = face_vertices_z.shape[0]
    b = first_idx_face_per_mesh.shape[0] - 1
    c = face_features.shape[2]
    assert face_vertices_z.shape == (n, 3) and face_vertices_image.shape == (n, 3, 2)
    assert face_bboxes.shape == (n, 4) and face_features.shape == (n, 3, c)
    opt = dict(dtype=face_vertices_z.dtype)
    idx = torch.full((b, height, width), -1, dtype=torch.long)
    w = torch.zeros((b, height, width, 3), **opt)
    feat = torch.zeros((b, height, width, c), **opt)
    r0, r1 = (0, -1) if rows is None else rows
    fn = getattr(lib(), "oracle_packed_rasterize_forward_" + _suf(face_vertices_z))
    fn(_p(face_vertices_z), _p(face_vertices_image), _p(face_bboxes), _p(face_features),
       _p(first_idx_face_per_mesh), _p(idx), _p(w), _p(feat), b, height, width, c,
       ctypes.c_float(multiplier), ctypes.c_float(eps), r0, r1)
    return [feat, idx, w]


def rasterize_backward_cuda(grad_interpolated_features, interpolated_features, selected_face_idx,
                            output_weights, face_vertices_image, face_features, eps, rows=None):
    """rasterization.cpp:106-168 and rasterization_cuda.cu:238-402."""
    _chk(grad_interpolated_features, selected_face_idx, output_weights, face_vertices_image,
         face_features)
    b, h, w, c = grad_interpolated_features.shape
    f = face_vertices_image.shape[1]
    gi = torch.zeros_like(face_vertices_image)
    gf = torch.zeros_like(face_features)
    r0, r1 = (0, -1) if rows is None else rows
    fn = getattr(lib(), "oracle_rasterize_backward_" + _suf(face_vertices_image))
    fn(_p(grad_interpolated_features), _p(selected_face_idx), _p(output_weights),
       _p(face_vertices_image), _p(face_features), _p(gi), _p(gf), b, h, w, f, c,
       ctypes.c_float(eps), r0, r1)
    return [gi, gf]


def dibr_soft_mask_forward_cuda(face_vertices_image, face_large_bboxes, selected_face_idx, sigmainv,
                                knum, multiplier, rows=None, side_buffers=True):
    """dibr_soft_mask.cpp:48-108 and dibr_soft_mask_cuda.cu:27-184."""
    _chk(face_vertices_image, face_large_bboxes, selected_face_idx)
    b, f = face_vertices_image.shape[:2]
    h, w = selected_face_idx.shape[1:]
    assert face_large_bboxes.shape == (b, f, 4)
    opt = dict(dtype=face_vertices_image.dtype)
    soft = torch.zeros((b, h, w), **opt)
    if side_buffers:
        cidx = torch.full((b, h, w, knum), -1, dtype=torch.long)
        prob = torch.zeros((b, h, w, knum), **opt)
        ctype = torch.zeros((b, h, w, knum), dtype=torch.uint8)
    else:
        cidx = prob = ctype = None
    r0, r1 = (0, -1) if rows is None else rows
    fn = getattr(lib(), "oracle_dibr_soft_mask_forward_" + _suf(face_vertices_image))
    fn(_p(face_vertices_image), _p(face_large_bboxes), _p(selected_face_idx), _p(prob), _p(cidx),
       _p(ctype), _p(soft), b, h, w, f, int(knum), ctypes.c_float(sigmainv),
       ctypes.c_float(multiplier), r0, r1)
    return [soft, prob, cidx, ctype]


def dibr_soft_mask_backward_cuda(grad_soft_mask, soft_mask, selected_face_idx, close_face_prob,
                                 close_face_idx, close_face_dist_type, face_vertices_image, sigmainv,
                                 multiplier, rows=None):
    """dibr_soft_mask.cpp:110-183 and dibr_soft_mask_cuda.cu:230-353."""
    _chk(grad_soft_mask, soft_mask, selected_face_idx, close_face_prob, close_face_idx,
         close_face_dist_type, face_vertices_image)
    b, f = face_vertices_image.shape[:2]
    h, w = selected_face_idx.shape[1:]
    knum = close_face_idx.shape[-1]
    gi = torch.zeros_like(face_vertices_image)
    r0, r1 = (0, -1) if rows is None else rows
    fn = getattr(lib(), "oracle_dibr_soft_mask_backward_" + _suf(face_vertices_image))
    fn(_p(grad_soft_mask), _p(soft_mask), _p(selected_face_idx), _p(close_face_prob),
       _p(close_face_idx), _p(close_face_dist_type), _p(face_vertices_image), _p(gi), b, h, w, f,
       knum, ctypes.c_float(sigmainv), ctypes.c_float(multiplier), r0, r1)
    return gi


def deftet_sparse_render_forward_cuda(face_vertices_z, face_vertices_image, face_bboxes, pixel_coords,
                                      pixel_depth_ranges, knum, eps):
    """deftet.cpp:47-109 (checks + allocation) and deftet_cuda.cu:31-190."""
    _chk(face_vertices_z, face_vertices_image, face_bboxes, pixel_coords, pixel_depth_ranges)
    b, f = face_vertices_z.shape[:2]
    p = pixel_coords.shape[1]
    assert face_vertices_image.shape == (b, f, 3, 2) and face_bboxes.shape == (b, f, 4)
    assert pixel_coords.shape == (b, p, 2) and pixel_depth_ranges.shape == (b, p, 2)
    opt = dict(dtype=face_vertices_z.dtype)
    idx = torch.full((b, p, knum), -1, dtype=torch.long)
    depth = torch.full((b, p, knum), -float("inf"), **opt)
    w0 = torch.zeros((b, p, knum), **opt)
    w1 = torch.zeros((b, p, knum), **opt)
    fn = getattr(lib(), "oracle_deftet_forward_" + _suf(face_vertices_z))
    fn(_p(face_vertices_z), _p(face_vertices_image), _p(face_bboxes), _p(pixel_coords), _p(pixel_depth_ranges),
       _p(idx), _p(depth), _p(w0), _p(w1), b, f, p, int(knum), ctypes.c_float(eps))
    return [idx, depth, w0, w1]


def deftet_sparse_render_backward_cuda(grad_interpolated_features, face_idx, weights, face_vertices_image,
                                       face_features, eps):
    """deftet.cpp:111-170 and deftet_cuda.cu:240-410."""
    _chk(grad_interpolated_features, face_idx, weights, face_vertices_image, face_features)
    b, p, knum, c = grad_interpolated_features.shape
    f = face_vertices_image.shape[1]
    gi = torch.zeros_like(face_vertices_image)
    gf = torch.zeros_like(face_features)
    fn = getattr(lib(), "oracle_deftet_backward_" + _suf(face_vertices_image))
    fn(_p(grad_interpolated_features), _p(face_idx), _p(weights), _p(face_vertices_image), _p(face_features),
       _p(gi), _p(gf), b, f, p, knum, c, ctypes.c_float(eps))
    return [gi, gf]


_mesh = types.SimpleNamespace(
    deftet_sparse_render_forward_cuda=deftet_sparse_render_forward_cuda,
    deftet_sparse_render_backward_cuda=deftet_sparse_render_backward_cuda,
    packed_rasterize_forward_cuda=packed_rasterize_forward_cuda,
    rasterize_backward_cuda=rasterize_backward_cuda,
    dibr_soft_mask_forward_cuda=dibr_soft_mask_forward_cuda,
    dibr_soft_mask_backward_cuda=dibr_soft_mask_backward_cuda,
)
OracleC = types.SimpleNamespace(render=types.SimpleNamespace(mesh=_mesh))


# ------------------------------------------------------------------ torch glue restated on CPU
def load_ref_C():
    """The reference's own CUDA ops built by oracle/build.py (oracle/_ref/ref_C.so); None if absent."""
    import importlib.util
    path = os.path.join(_HERE, "_ref", "ref_C.so")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("ref_C", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def rasterize_forward(height, width, face_vertices_z, face_vertices_image, face_features,
                      valid_faces=None, multiplier=None, eps=None, rows=None, backend=None):
    """RasterizeCuda.forward, rasterization.py:290-369 (returns features, face_idx, weights).
    ``backend`` = an object with ``render.mesh.packed_rasterize_forward_cuda`` (default: the C++ oracle;
    pass ``load_ref_C()`` with CUDA tensors to run the reference's own kernels through the same glue)."""
    dev = face_vertices_z.device
    multiplier = 1000 if multiplier is None else multiplier       # rasterization.py:469-475
    eps = 1e-8 if eps is None else eps
    if isinstance(face_features, (list, tuple)):
        face_features = torch.cat(face_features, dim=-1)           # :480-481
    b, f = face_vertices_z.shape[:2]
    c = face_features.shape[-1]
    face_features = face_features.contiguous()
    face_vertices_image = face_vertices_image.contiguous()
    if valid_faces is None:                                        # :309-323
        vi = (torch.arange(b, device=dev).reshape(-1, 1).repeat(1, f).reshape(-1),
              torch.arange(f, device=dev).reshape(1, -1).repeat(b, 1).reshape(-1))
        v_img = face_vertices_image.reshape(b * f, 3, 2)
        v_z = face_vertices_z.reshape(b * f, 3)
        v_feat = face_features.reshape(b * f, 3, c)
        nfm = torch.full((b,), f, dtype=torch.long, device=dev)
    else:                                                          # :324-332
        vi = torch.where(valid_faces)
        v_img = face_vertices_image[vi[0], vi[1]]
        v_z = face_vertices_z[vi[0], vi[1]]
        v_feat = face_features[vi[0], vi[1]]
        nfm = torch.sum(valid_faces.reshape(b, -1), dim=1)
    first = torch.zeros(b + 1, dtype=torch.long, device=dev)       # :333-334
    torch.cumsum(nfm, dim=0, out=first[1:])
    v_img = v_img * multiplier                                     # :337
    pmin = torch.min(v_img, dim=1)[0]                              # :342-344
    pmax = torch.max(v_img, dim=1)[0]
    bboxes = torch.cat((pmin, pmax), dim=1)
    if backend is None:
        feat, sel, w = packed_rasterize_forward_cuda(
            height, width, v_z.contiguous(), v_img.contiguous(), bboxes.contiguous(),
            v_feat.contiguous(), first.contiguous(), multiplier, eps, rows=rows)
    else:
        feat, sel, w = backend.render.mesh.packed_rasterize_forward_cuda(
            height, width, v_z.contiguous(), v_img.contiguous(), bboxes.contiguous(),
            v_feat.contiguous(), first.contiguous(), multiplier, eps)
    if vi[1].numel() > 0:                                          # :357-363
        face_idx = vi[1][(sel + first[:-1].reshape(-1, 1, 1)).reshape(-1).clamp(min=0, max=vi[1].numel() - 1)]
        face_idx = face_idx.reshape(sel.shape).contiguous()
    else:
        face_idx = torch.full_like(sel, -1)
    face_idx[sel == -1] = -1
    return feat, face_idx, w


def rasterize_backward(grad_features, face_idx, weights, face_vertices_image, face_features, eps=None,
                       rows=None, backend=None, interpolated_features=None):
    """RasterizeCuda.backward, rasterization.py:371-388."""
    eps = 1e-8 if eps is None else eps
    if isinstance(face_features, (list, tuple)):
        face_features = torch.cat(face_features, dim=-1)
    if backend is not None:
        return backend.render.mesh.rasterize_backward_cuda(
            grad_features.contiguous(), interpolated_features, face_idx, weights,
            face_vertices_image.contiguous(), face_features.contiguous(), eps)
    return rasterize_backward_cuda(grad_features.contiguous(), None, face_idx, weights,
                                   face_vertices_image.contiguous(), face_features.contiguous(), eps,
                                   rows=rows)


def soft_mask_bboxes(face_vertices_image, boxlen, multiplier):
    """dibr.py:31-39: multiplied coordinates and enlarged bboxes."""
    img = face_vertices_image.contiguous() * multiplier
    pmin = torch.min(img, dim=-2)[0]
    pmax = torch.max(img, dim=-2)[0]
    bb = torch.cat([pmin - boxlen * multiplier, pmax + boxlen * multiplier], dim=-1)
    return img, bb.contiguous()


def dibr_soft_mask_forward(face_vertices_image, selected_face_idx, sigmainv=7000, boxlen=0.02, knum=30,
                           multiplier=1000., rows=None, side_buffers=True, backend=None):
    """DibrSoftMaskCuda.forward, dibr.py:29-55.  Returns (soft, multiplied img, prob, idx, type)."""
    img, bb = soft_mask_bboxes(face_vertices_image, boxlen, multiplier)
    if backend is not None:
        soft, prob, cidx, ctype = backend.render.mesh.dibr_soft_mask_forward_cuda(
            img, bb, selected_face_idx.contiguous(), sigmainv, knum, multiplier)
        return soft, img, prob, cidx, ctype
    soft, prob, cidx, ctype = dibr_soft_mask_forward_cuda(
        img, bb, selected_face_idx.contiguous(), sigmainv, knum, multiplier, rows=rows,
        side_buffers=side_buffers)
    return soft, img, prob, cidx, ctype


def dibr_soft_mask_backward(grad_soft, soft, img_mult, selected_face_idx, prob, cidx, ctype, sigmainv,
                            multiplier, rows=None, backend=None):
    """DibrSoftMaskCuda.backward, dibr.py:57-73."""
    if backend is not None:
        return backend.render.mesh.dibr_soft_mask_backward_cuda(
            grad_soft.contiguous(), soft, selected_face_idx, prob, cidx, ctype, img_mult, sigmainv, multiplier)
    return dibr_soft_mask_backward_cuda(grad_soft.contiguous(), soft, selected_face_idx, prob, cidx,
                                        ctype, img_mult, sigmainv, multiplier, rows=rows)


def dibr_rasterization_forward(height, width, face_vertices_z, face_vertices_image, face_features,
                               face_normals_z, sigmainv=7000, boxlen=0.02, knum=30, multiplier=None,
                               eps=None):
    """dibr_rasterization, dibr.py:119-209 (forward values only)."""
    feat, face_idx, w = rasterize_forward(height, width, face_vertices_z, face_vertices_image,
                                          face_features, face_normals_z >= 0., multiplier, eps)
    m = 1000. if multiplier is None else multiplier
    soft = dibr_soft_mask_forward(face_vertices_image, face_idx, sigmainv, boxlen, knum, m)[0]
    return feat, soft, face_idx


# ---------------------------------------------------------------------------------------------
# prepare_vertices — CPU restatement in torch ops, differentiable through autograd.
# Parity status: pinned against tests/golden/prepare_vertices.npz (outputs and gradients produced by the
# reference's own modules, tests/golden/make_golden_prepare.py).
def rotate_translate_points(points, camera_rot, camera_trans):
    """kaolin/render/camera/legacy.py:22-37: (points - trans) @ rot^T."""
    translated = points - camera_trans.view(-1, 1, 3)
    return torch.matmul(translated, camera_rot.permute(0, 2, 1))


def perspective_camera(points, camera_proj):
    """kaolin/render/camera/legacy.py:120-139: (x*p0, y*p1) / (z*p2)."""
    projected = points * camera_proj.view(-1, 1, 3)
    return projected[:, :, :2] / projected[:, :, 2:3]


def index_vertices_by_faces(vertices_features, faces):
    """kaolin/ops/mesh/mesh.py:24-45: gather (B,V,k) by (F,3) -> (B,F,3,k)."""
    return vertices_features[:, faces]


def face_normals(face_vertices, unit=False):
    """kaolin/ops/mesh/trianglemesh.py:313-336."""
    e0 = face_vertices[:, :, 1] - face_vertices[:, :, 0]
    e1 = face_vertices[:, :, 2] - face_vertices[:, :, 0]
    n = torch.cross(e0, e1, dim=2)
    if unit:
        n = n / (n.norm(dim=2, keepdim=True) + 1e-10)
    return n


def prepare_vertices(vertices, faces, camera_proj, camera_rot=None, camera_trans=None, camera_transform=None):
    """kaolin/render/mesh/utils.py:128-175."""
    if camera_transform is None:
        assert camera_trans is not None and camera_rot is not None
        vertices_camera = rotate_translate_points(vertices, camera_rot, camera_trans)
    else:
        assert camera_trans is None and camera_rot is None
        padded = torch.nn.functional.pad(vertices, (0, 1), mode='constant', value=1.)
        vertices_camera = padded @ camera_transform
    vertices_image = perspective_camera(vertices_camera, camera_proj)
    face_vertices_camera = index_vertices_by_faces(vertices_camera, faces)
    face_vertices_image = index_vertices_by_faces(vertices_image, faces)
    return face_vertices_camera, face_vertices_image, face_normals(face_vertices_camera, unit=True)


# ---------------------------------------------------------------------------------------------
# texture_mapping and mask_iou (row 3) — CPU restatements in torch ops.
# Parity status: pinned against the reference's known-answer tests
# (tests/python/kaolin/render/mesh/test_utils.py:61-112, tests/python/kaolin/metrics/test_render.py:24-53),
# restated in tests/test_texture_losses.py.
def texture_mapping(texture_coordinates, texture_maps, mode='nearest'):
    """kaolin/render/mesh/utils.py:23-76."""
    batch_size = texture_coordinates.shape[0]
    num_channels = texture_maps.shape[1]
    coords = texture_coordinates.reshape(batch_size, -1, 1, 2)
    coords = torch.clamp(coords, 0., 1.) * 2 - 1                       # utils.py:62-63
    coords = torch.stack([coords[..., 0], -coords[..., 1]], dim=-1)    # utils.py:64 (reverse y), out of place
    sampled = torch.nn.functional.grid_sample(texture_maps, coords, mode=mode, align_corners=False,
                                              padding_mode='border')
    sampled = sampled.permute(0, 2, 3, 1)
    return sampled.reshape(batch_size, *texture_coordinates.shape[1:-1], num_channels)


def mask_iou(lhs_mask, rhs_mask):
    """kaolin/metrics/render.py:18-40."""
    batch_size = lhs_mask.shape[0]
    assert rhs_mask.shape == lhs_mask.shape
    sil_mul = lhs_mask * rhs_mask
    sil_add = lhs_mask + rhs_mask
    iou_up = torch.sum(sil_mul.reshape(batch_size, -1), dim=1)
    iou_down = torch.sum((sil_add - sil_mul).reshape(batch_size, -1), dim=1)
    return 1.0 - torch.mean(iou_up / (iou_down + 1e-10))



# ---------------------------------------------------------------------------------------------
# deftet_sparse_render — torch glue of DeftetSparseRenderer restated on top of the
# native-op-shaped functions above (backend=None: the C++ oracle on CPU tensors; backend=ref_C: the
# reference's own CUDA kernels on GPU tensors).
def deftet_sparse_render_forward(pixel_coords, render_ranges, face_vertices_z, face_vertices_image,
                                 face_features, knum=300, eps=1e-8, backend=None):
    """DeftetSparseRenderer.forward, deftet.py:272-318.  Returns (interpolated_features, sorted_face_idx,
    weights).  The depth sort is made stable here (ties keep insertion = face order); torch.argsort of the
    reference leaves the order of equal depths unspecified."""
    b, f = face_vertices_z.shape[:2]
    c = face_features.shape[-1]
    p = pixel_coords.shape[1]
    face_min = torch.min(face_vertices_image, dim=2)[0]
    face_max = torch.max(face_vertices_image, dim=2)[0]
    bboxes = torch.cat((face_min, face_max), dim=2)
    mesh = OracleC.render.mesh if backend is None else backend.render.mesh
    face_idx, depth, w0, w1 = mesh.deftet_sparse_render_forward_cuda(
        face_vertices_z.contiguous(), face_vertices_image.contiguous(), bboxes.contiguous(),
        pixel_coords.contiguous(), render_ranges.contiguous(), knum, eps)
    order = torch.sort(depth, descending=True, dim=-1, stable=True)[1]
    s_idx = torch.gather(face_idx, -1, order).contiguous()
    s_w0 = torch.gather(w0, -1, order)
    s_w1 = torch.gather(w1, -1, order)
    s_w2 = (s_idx != -1).float() - (s_w0 + s_w1)
    gidx = (s_idx + 1).reshape(b, -1, 1, 1).expand(b, p * knum, 3, c)
    sel = torch.gather(torch.nn.functional.pad(face_features, (0, 0, 0, 0, 1, 0), value=0.), 1, gidx)
    sel = sel.reshape(b, p, knum, 3, c)
    weights = torch.stack([s_w0, s_w1, s_w2], dim=-1).contiguous()
    feats = torch.sum(weights.unsqueeze(-1) * sel, dim=-2).contiguous()
    return feats, s_idx, weights


def deftet_sparse_render_backward(grad_features, face_idx, weights, face_vertices_image, face_features,
                                  eps=1e-8, backend=None):
    """DeftetSparseRenderer.backward, deftet.py:320-337."""
    mesh = OracleC.render.mesh if backend is None else backend.render.mesh
    return mesh.deftet_sparse_render_backward_cuda(grad_features.contiguous(), face_idx, weights,
                                                   face_vertices_image.contiguous(), face_features.contiguous(),
                                                   eps)
