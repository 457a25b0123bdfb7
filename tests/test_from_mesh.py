"""dibr_rasterization_from_mesh — prepare_vertices folded into the setup kernel — against the
unfused chain ``prepare_vertices -> dibr_rasterization`` of the reference (kaolin/render/mesh/utils.py:128-175,
dibr.py:119-209; examples/tutorial/ian_dibr.py:246-252), which the other GPU suites pin to the reference's kernels.

Forward values must be IDENTICAL (the setup kernel projects with prepare_vertices' arithmetic, so the per-tile kernels
see the same bits): ``torch.equal`` on face_idx, features and soft mask.  Gradients are sums of the same terms in a
different order (unordered atomics on both sides): 1e-4 of the largest component, the tolerance `north_star` states.
"""
import math

import pytest
import torch

from helpers import norm_close

pytestmark = pytest.mark.gpu


def _cameras(B, dtype, dev, radius=3.0):
    from kaolin_b200 import synth
    rot, eye = synth.orbit_cameras(B, radius=radius, dtype=torch.float64)
    proj = torch.tensor([1 / math.tan(math.pi / 8), 1 / math.tan(math.pi / 8), -1.0], dtype=dtype, device=dev)
    return rot.to(dtype).to(dev), eye.to(dtype).to(dev), proj


def _sphere(n_lon, n_lat, dtype, dev, seed=5):
    from kaolin_b200 import synth
    v, f = synth.uv_sphere(n_lon, n_lat, 0.05, torch.Generator().manual_seed(seed))
    v = v / v.norm(dim=1).max()
    return v.to(dtype).to(dev), f.to(dev)


def _chain(H, W, verts, faces, feats, proj, rot=None, trans=None, transform=None, **kw):
    from kaolin_b200.render.mesh import prepare_vertices, dibr_rasterization
    cam, img, nrm = prepare_vertices(verts, faces, proj, rot, trans, transform)
    return dibr_rasterization(H, W, cam[..., 2], img, feats, nrm[..., 2], **kw)


def _fused(H, W, verts, faces, feats, proj, rot=None, trans=None, transform=None, **kw):
    from kaolin_b200.render.mesh import dibr_rasterization_from_mesh
    return dibr_rasterization_from_mesh(H, W, verts, faces, feats, proj, rot, trans, transform, **kw)


def _compare(H, W, verts, faces, feats, proj, rot=None, trans=None, transform=None, cam_grad=False, seed=0, **kw):
    g = torch.Generator().manual_seed(100 + seed)
    is_list = isinstance(feats, (list, tuple))
    flist = list(feats) if is_list else [feats]
    B = (rot if rot is not None else transform).shape[0]
    g_out = [torch.rand(B, H, W, t.shape[-1], generator=g, dtype=torch.float64).to(t) for t in flist]
    g_soft = torch.rand(B, H, W, generator=g, dtype=torch.float64).to(verts)
    res = []
    for fn in (_chain, _fused):
        leaf = verts.detach().clone().requires_grad_(True)
        fl = [t.detach().clone().requires_grad_(True) for t in flist]
        cams = {}
        for name, t in (("rot", rot), ("trans", trans), ("transform", transform)):
            cams[name] = None if t is None else t.detach().clone().requires_grad_(cam_grad)
        pr = proj.detach().clone().requires_grad_(cam_grad)
        out, soft, idx = fn(H, W, leaf, faces, fl if is_list else fl[0], pr, cams["rot"], cams["trans"],
                            cams["transform"], **kw)
        outs = list(out) if is_list else [out]
        torch.autograd.backward(outs + [soft], g_out + [g_soft])
        grads = {"vertices": leaf.grad}
        for i, t in enumerate(fl):
            grads[f"feat{i}"] = t.grad
        if cam_grad:
            grads["proj"] = pr.grad
            for name, t in cams.items():
                if t is not None:
                    grads[name] = t.grad
        res.append((outs, soft, idx, grads))
    (o0, s0, i0, g0), (o1, s1, i1, g1) = res
    assert torch.equal(i0, i1), "face_idx differs from the unfused chain"
    assert int((i0 >= 0).sum()) > 0 and int((i0 < 0).sum()) > 0, "scene must have covered and uncovered pixels"
    for a, b in zip(o0, o1):
        assert torch.equal(a, b), "interpolated features differ from the unfused chain"
    assert torch.equal(s0, s1), "soft mask differs from the unfused chain"
    for k in g0:
        assert g1[k] is not None, k
        norm_close(g1[k], g0[k], 1e-4, f"grad {k}")


@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("B", [1, 5])
def test_shared_mesh_rot_trans(dname, B):
    dtype = torch.float32 if dname == "f32" else torch.float64
    dev = "cuda"
    v, f = _sphere(40, 24, dtype, dev)
    rot, eye, proj = _cameras(B, dtype, dev)
    feats = torch.rand(B, f.shape[0], 3, 3, device=dev, dtype=dtype)
    _compare(96, 80, v.unsqueeze(0), f, feats, proj, rot, eye, seed=B)


def test_per_view_meshes_and_feature_list():
    dev, dtype, B = "cuda", torch.float32, 3
    v, f = _sphere(32, 20, dtype, dev)
    verts = v.unsqueeze(0) + 0.01 * torch.rand(B, v.shape[0], 3, device=dev)
    rot, eye, proj = _cameras(B, dtype, dev)
    feats = [torch.rand(B, f.shape[0], 3, c, device=dev) for c in (2, 3, 1)]
    _compare(64, 64, verts, f, feats, proj, rot, eye, seed=7)


@pytest.mark.parametrize("dname", ["f32", "f64"])
def test_camera_transform_form_and_camera_gradients(dname):
    dtype = torch.float32 if dname == "f32" else torch.float64
    dev, B = "cuda", 4
    v, f = _sphere(36, 18, dtype, dev)
    rot, eye, proj = _cameras(B, dtype, dev)
    # (p - t) R^T = p R^T - t R^T  ->  transform = [R^T ; -t R^T]   (utils.py:163-167)
    rt = rot.transpose(1, 2)
    transform = torch.cat([rt, -(eye.unsqueeze(1) @ rt)], dim=1).contiguous()
    feats = torch.rand(B, f.shape[0], 3, 3, device=dev, dtype=dtype)
    _compare(72, 72, v.unsqueeze(0), f, feats, proj, transform=transform, cam_grad=True, seed=3)
    _compare(72, 72, v.unsqueeze(0), f, feats, proj, rot, eye, cam_grad=True, seed=4)


def test_non_default_parameters_and_ragged_image():
    dev, dtype, B = "cuda", torch.float32, 2
    v, f = _sphere(30, 16, dtype, dev)
    rot, eye, proj = _cameras(B, dtype, dev)
    feats = torch.rand(B, f.shape[0], 3, 4, device=dev)
    _compare(35, 51, v.unsqueeze(0), f, feats, proj, rot, eye, seed=9, sigmainv=70., boxlen=0.05, knum=10,
             multiplier=100., eps=1e-6)


def test_c4_shape_two_cameras_full_size():
    """The benchmarked c4 call (40k-face sphere, 512 x 512), two cameras."""
    from kaolin_b200 import synth
    dev = "cuda"
    d = synth.make_inputs("c4", views=[0, 37])
    v, f = d["vertices"].to(dev), d["faces"].to(dev)
    proj = torch.tensor([1 / math.tan(math.pi / 8), 1 / math.tan(math.pi / 8), -1.0], device=dev)
    feats = d["face_features"][0].to(dev)
    _compare(d["H"], d["W"], v.unsqueeze(0), f, feats, proj, d["rot"].to(dev), d["eye"].to(dev), seed=11)


def test_only_features_require_grad_and_no_grad():
    from kaolin_b200.render.mesh import dibr_rasterization_from_mesh
    dev, B = "cuda", 2
    v, f = _sphere(24, 12, torch.float32, dev)
    rot, eye, proj = _cameras(B, torch.float32, dev)
    feats = torch.rand(B, f.shape[0], 3, 3, device=dev, requires_grad=True)
    out, soft, idx = dibr_rasterization_from_mesh(48, 48, v.unsqueeze(0), f, feats, proj, rot, eye)
    out.sum().backward()
    assert feats.grad is not None and torch.isfinite(feats.grad).all() and float(feats.grad.abs().sum()) > 0
    with torch.no_grad():
        out2, soft2, idx2 = dibr_rasterization_from_mesh(48, 48, v.unsqueeze(0), f, feats.detach(), proj, rot, eye)
    assert torch.equal(idx, idx2) and torch.equal(out.detach(), out2) and torch.equal(soft.detach(), soft2)


def test_errors():
    from kaolin_b200.render.mesh import dibr_rasterization_from_mesh
    dev = "cuda"
    v, f = _sphere(12, 8, torch.float32, dev)
    rot, eye, proj = _cameras(2, torch.float32, dev)
    feats = torch.rand(2, f.shape[0], 3, 3, device=dev)
    with pytest.raises(AssertionError):
        dibr_rasterization_from_mesh(32, 32, v.unsqueeze(0), f, feats, proj)                 # no camera
    with pytest.raises(RuntimeError):
        dibr_rasterization_from_mesh(32, 32, v.unsqueeze(0), f, feats[:1], proj, rot, eye)   # feature batch mismatch
    with pytest.raises(RuntimeError):
        dibr_rasterization_from_mesh(32, 32, v.unsqueeze(0).cpu(), f, feats, proj, rot, eye)
    with pytest.raises(ValueError):
        dibr_rasterization_from_mesh(32, 32, v.unsqueeze(0), f, [], proj, rot, eye)
