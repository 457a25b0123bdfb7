"""prepare_vertices (kaolin/render/mesh/utils.py:128-175).

CPU: the oracle restatement against golden vectors produced by the reference's own modules
(tests/golden/make_golden_prepare.py -> tests/golden/prepare_vertices.npz), outputs and gradients.
GPU: the fused CUDA kernels (through the C ABI) against the same golden vectors and against the oracle on
seeded inputs.  Tolerance: 1e-5 relative for fp32 (BASELINE.json north_star: features within 1e-5), 1e-12 for
fp64 — floating point, no bit-exactness claim (the reference's matmul has no fixed summation order).
Unit normals of small faces are ill-conditioned in fp32 (cross product of short edges of coordinates of
magnitude ~3): the reference's OWN two fp32 formulations (rot/trans vs 4x3 transform) differ by 1.3e-5 on the
golden mesh's normals and 1.4e-5 on its gradients, so fp32 normals and gradients are held to 1e-4 (the
north_star gradient tolerance) and that spread is asserted below."""
import math

import numpy as np
import pytest
import torch

from helpers import golden
from oracle import oracle as O

TOL = {"f32": 1e-5, "f64": 1e-12}
TOL_ILL = {"f32": 1e-4, "f64": 1e-10}          # normals / gradients (see the module docstring)
DT = {"f32": torch.float32, "f64": torch.float64}


def rel_close(a, b, tol):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    scale = max(float(b.abs().max()), 1e-30)
    return float((a - b).abs().max()) <= tol * scale


def _case(dname, mode, layout, device="cpu"):
    g = golden("prepare_vertices")
    t = lambda k: torch.from_numpy(g[f"{dname}_{k}"]).to(device)  # noqa: E731
    verts = t("vertices_b") if layout == "b" else t("vertices_shared").unsqueeze(0)
    cams = dict(camera_rot=t("rot"), camera_trans=t("trans")) if mode == "rt" else dict(camera_transform=t("transform"))
    want = {k: t(f"{mode}_{layout}_{k}") for k in ("cam", "img", "nrm", "grad_vertices")}
    grads = [t("g_cam"), t("g_img"), t("g_nrm")]
    return verts, torch.from_numpy(g["faces"]).to(device), t("proj"), cams, grads, want


CASES = [(d, m, l) for d in ("f32", "f64") for m in ("rt", "tf") for l in ("b", "shared")]


@pytest.mark.parametrize("dname,mode,layout", CASES)
def test_oracle_matches_reference_golden(dname, mode, layout):
    verts, faces, proj, cams, grads, want = _case(dname, mode, layout)
    leaf = verts.clone().requires_grad_(True)
    src = leaf if layout == "b" else leaf.expand(3, -1, -1)
    out = O.prepare_vertices(src, faces, proj, **cams)
    torch.autograd.backward(list(out), grads)
    tol = TOL[dname]
    for o, k in zip(out, ("cam", "img", "nrm")):
        assert o.shape == want[k].shape
        assert rel_close(o, want[k], TOL_ILL[dname] if k == "nrm" else tol), k
    assert rel_close(leaf.grad.reshape(want["grad_vertices"].shape), want["grad_vertices"], TOL_ILL[dname])


def test_golden_fixture_is_self_consistent():
    g = golden("prepare_vertices")
    assert g["faces"].dtype == np.int64 and g["faces"].shape[1] == 3
    assert g["f32_rt_b_cam"].shape == (3, g["faces"].shape[0], 3, 3)
    # rot/trans and the 4x3 transform describe the same cameras in the reference
    assert np.allclose(g["f64_rt_b_cam"], g["f64_tf_b_cam"], rtol=1e-9, atol=1e-12)
    n = g["f64_rt_b_nrm"]
    assert np.allclose(np.linalg.norm(n, axis=-1), 1.0, atol=1e-8)
    # the reference's own fp32 spread between its two formulations (why TOL_ILL is 1e-4 and not 1e-5)
    for k, lo, hi in (("nrm", 1e-6, 5e-5), ("grad_vertices", 1e-6, 5e-5), ("cam", 0, 1e-6), ("img", 0, 2e-6)):
        a, b = g["f32_rt_shared_" + k].astype(np.float64), g["f32_tf_shared_" + k].astype(np.float64)
        spread = np.abs(a - b).max() / np.abs(a).max()
        assert lo <= spread <= hi, (k, spread)


# ----------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("dname,mode,layout", CASES)
def test_cuda_matches_reference_golden(dname, mode, layout):
    from kaolin_b200.render.mesh import prepare_vertices
    verts, faces, proj, cams, grads, want = _case(dname, mode, layout, "cuda")
    leaf = verts.clone().requires_grad_(True)
    out = prepare_vertices(leaf, faces, proj, **cams)
    torch.autograd.backward(list(out), grads)
    tol = TOL[dname]
    for o, k in zip(out, ("cam", "img", "nrm")):
        assert o.shape == want[k].shape and o.is_contiguous()
        assert rel_close(o, want[k], TOL_ILL[dname] if k == "nrm" else tol), k
    assert rel_close(leaf.grad.reshape(want["grad_vertices"].shape), want["grad_vertices"], TOL_ILL[dname])


def _random_scene(dtype, B, V, F, seed, device):
    g = torch.Generator().manual_seed(seed)
    verts = (torch.rand(1, V, 3, generator=g, dtype=torch.float64) - 0.5)
    if B > 1:                                        # per-view meshes: one shape, jittered per view
        verts = verts + 0.01 * torch.rand(B, V, 3, generator=g, dtype=torch.float64)
    verts = verts.to(dtype)
    faces = torch.randint(0, V, (F, 3), generator=g)
    # keep well-conditioned triangles only (fp32 unit normals of slivers are noise in the reference as well):
    # sin of the angle at corner 0 above 0.2 and edges longer than 0.1, judged on the first mesh of the batch
    e0 = verts[0, faces[:, 1]].double() - verts[0, faces[:, 0]].double()
    e1 = verts[0, faces[:, 2]].double() - verts[0, faces[:, 0]].double()
    sin = torch.cross(e0, e1, dim=1).norm(dim=1) / (e0.norm(dim=1) * e1.norm(dim=1) + 1e-30)
    faces = faces[(sin > 0.2) & (e0.norm(dim=1) > 0.1) & (e1.norm(dim=1) > 0.1)]
    ang = torch.rand(B, generator=g, dtype=torch.float64) * 2 * math.pi
    eye = torch.stack([3 * torch.sin(ang), torch.full_like(ang, 0.7), 3 * torch.cos(ang)], -1)
    cz = -eye / eye.norm(dim=1, keepdim=True)
    cx = torch.cross(cz, torch.tensor([[0., 1., 0.]], dtype=torch.float64).expand_as(cz), dim=1)
    cx = cx / cx.norm(dim=1, keepdim=True)
    cy = torch.cross(cx, cz, dim=1)
    rot = torch.stack([cx, cy, -cz], 1).to(dtype)
    proj = torch.tensor([[2.4142], [2.0], [-1.0]], dtype=dtype)
    return [t.to(device) for t in (verts, faces, rot, eye.to(dtype), proj)]


@pytest.mark.gpu
@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("shared", [False, True])
@pytest.mark.parametrize("which", ["all", "img", "nrm", "cam"])
def test_cuda_matches_oracle_random(dname, shared, which):
    """Seeded random meshes (including degenerate faces with repeated vertices), each upstream gradient alone
    and all together; unused outputs contribute nothing."""
    from kaolin_b200.render.mesh import prepare_vertices
    dtype = DT[dname]
    B, V, F = 5, 700, 3001
    verts, faces, rot, trans, proj = _random_scene(dtype, B, V, F, 7, "cuda")
    if shared:
        verts = verts[:1].contiguous()
    faces[:8, 1] = faces[:8, 0]                      # degenerate triangles: zero normal, 0/(0+1e-10)
    la = verts.clone().requires_grad_(True)
    lb = verts.clone().requires_grad_(True)
    got = prepare_vertices(la, faces, proj, rot, trans)
    ref = O.prepare_vertices(lb.double().expand(B, -1, -1), faces, proj.double(), rot.double(), trans.double())
    tol = TOL[dname]
    for i, (a, b) in enumerate(zip(got, ref)):
        assert a.dtype == dtype and rel_close(a, b, TOL_ILL[dname] if i == 2 else tol)
    g = torch.Generator().manual_seed(11)
    ups = [torch.randn(o.shape, generator=g, dtype=torch.float64).to("cuda") for o in ref]
    # a zero-area face turns an upstream normal gradient into +-g/1e-10 on one and the same vertex (exact
    # cancellation in real arithmetic, rounding noise of order 1e3 in fp32 — in the reference too): no
    # upstream normal gradient on those eight faces
    ups[2][:, :8] = 0
    sel = {"all": (0, 1, 2), "cam": (0,), "img": (1,), "nrm": (2,)}[which]
    torch.autograd.backward([got[i] for i in sel], [ups[i].to(dtype) for i in sel])
    torch.autograd.backward([ref[i] for i in sel], [ups[i] for i in sel])
    assert rel_close(la.grad, lb.grad, TOL_ILL[dname])


@pytest.mark.gpu
@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("mode", ["rt", "tf"])
@pytest.mark.parametrize("shared", [False, True])
def test_cuda_camera_gradients_match_torch(dname, mode, shared):
    """The reference differentiates through camera_proj / camera_rot / camera_trans / camera_transform (pure torch,
    utils.py:160-172): the backward kernel returns those gradients too (warp-reduced sums), with or without a
    gradient for the vertices.  F is not a multiple of 32, so warps straddle two views."""
    from kaolin_b200.render.mesh import prepare_vertices
    dtype = DT[dname]
    B, V, F = 3, 400, 1003
    verts, faces, rot, trans, proj = _random_scene(dtype, B, V, F, 21, "cuda")
    if shared:
        verts = verts[:1].contiguous()
    m = torch.cat([rot.permute(0, 2, 1), -(trans.unsqueeze(1) @ rot.permute(0, 2, 1))], dim=1).contiguous()
    for verts_grad in (True, False):
        la = verts.clone().requires_grad_(verts_grad)
        lb = verts.double().clone().requires_grad_(verts_grad)
        pa = proj.clone().requires_grad_(True)
        pb = proj.double().clone().requires_grad_(True)
        if mode == "rt":
            ca = [rot.clone().requires_grad_(True), trans.clone().requires_grad_(True)]
            cb = [t.detach().double().clone().requires_grad_(True) for t in ca]
            got = prepare_vertices(la, faces, pa, ca[0], ca[1])
            ref = O.prepare_vertices(lb.expand(B, -1, -1), faces, pb, cb[0], cb[1])
        else:
            ca = [m.clone().requires_grad_(True)]
            cb = [m.double().clone().requires_grad_(True)]
            got = prepare_vertices(la, faces, pa, camera_transform=ca[0])
            ref = O.prepare_vertices(lb.expand(B, -1, -1), faces, pb, camera_transform=cb[0])
        g = torch.Generator().manual_seed(5)
        ups = [torch.randn(o.shape, generator=g, dtype=torch.float64).to("cuda") for o in ref]
        torch.autograd.backward(list(got), [u.to(dtype) for u in ups])
        torch.autograd.backward(list(ref), ups)
        tol = TOL_ILL[dname]
        assert rel_close(pa.grad, pb.grad, tol), (pa.grad, pb.grad)
        for a, b in zip(ca, cb):
            assert a.grad.shape == a.shape and rel_close(a.grad, b.grad, tol)
        if verts_grad:
            assert rel_close(la.grad, lb.grad, tol)
        else:
            assert la.grad is None


@pytest.mark.gpu
def test_cuda_transform_equals_rot_trans():
    from kaolin_b200.render.mesh import prepare_vertices
    verts, faces, rot, trans, proj = _random_scene(torch.float64, 4, 300, 900, 3, "cuda")
    # [v,1] @ M with M = [rot^T ; -trans @ rot^T]  (kaolin/render/camera/legacy.py:85-118)
    m = torch.cat([rot.permute(0, 2, 1), -(trans.unsqueeze(1) @ rot.permute(0, 2, 1))], dim=1)
    a = prepare_vertices(verts, faces, proj, rot, trans)
    b = prepare_vertices(verts, faces, proj, camera_transform=m)
    for x, y in zip(a, b):
        assert rel_close(x, y, 1e-12)


@pytest.mark.gpu
def test_cuda_empty_and_errors():
    from kaolin_b200.render.mesh import prepare_vertices
    verts, faces, rot, trans, proj = _random_scene(torch.float32, 2, 50, 20, 5, "cuda")
    out = prepare_vertices(verts, faces[:0], proj, rot, trans)
    assert [tuple(o.shape) for o in out] == [(2, 0, 3, 3), (2, 0, 3, 2), (2, 0, 3)]
    with pytest.raises(AssertionError):
        prepare_vertices(verts, faces, proj)
    with pytest.raises(AssertionError):
        prepare_vertices(verts, faces, proj, rot, trans, camera_transform=torch.zeros(2, 4, 3, device="cuda"))
    with pytest.raises(NotImplementedError):
        prepare_vertices(verts, torch.zeros(4, 4, dtype=torch.long, device="cuda"), proj, rot, trans)
    with pytest.raises(RuntimeError):
        prepare_vertices(verts.cpu(), faces.cpu(), proj.cpu(), rot.cpu(), trans.cpu())
    with pytest.raises(RuntimeError):
        prepare_vertices(verts.half(), faces, proj.half(), rot.half(), trans.half())
    bad = faces.clone()
    bad[3, 1] = 50                                   # out of range: NaN for that face only, no fault
    out = prepare_vertices(verts, bad, proj, rot, trans)
    assert torch.isnan(out[0][:, 3]).any() and not torch.isnan(out[0][:, :3]).any()
    torch.cuda.synchronize()


@pytest.mark.gpu
def test_cuda_feeds_dibr_rasterization_like_the_reference_pipeline():
    """prepare_vertices -> dibr_rasterization end to end: gradient w.r.t. the vertices equals the one obtained
    by running the oracle's torch prepare_vertices (autograd) in front of the same CUDA rasterizer."""
    from kaolin_b200 import synth
    from kaolin_b200.render.mesh import prepare_vertices, dibr_rasterization
    v, f = synth.icosphere(3)
    rot, eye = synth.orbit_cameras(4)
    dev = "cuda"
    # fp64: the two producers differ by rounding only, which must not flip a pixel's face at an edge
    v, f, rot, eye = v.double().to(dev), f.to(dev), rot.double().to(dev), eye.double().to(dev)
    proj = torch.tensor([1 / math.tan(math.pi / 8), 1 / math.tan(math.pi / 8), -1.0], device=dev, dtype=torch.float64)
    feats = torch.rand(1, f.shape[0], 3, 3, device=dev, dtype=torch.float64).expand(4, -1, -1, -1).contiguous()
    grads = []
    for prep in (prepare_vertices, O.prepare_vertices):
        leaf = v.clone().unsqueeze(0).requires_grad_(True)
        src = leaf if prep is prepare_vertices else leaf.expand(4, -1, -1)
        fcam, fimg, nrm = prep(src, f, proj, rot, eye)
        img, soft, idx = dibr_rasterization(64, 64, fcam[..., 2], fimg, feats, nrm[..., 2])
        (img.sum() + soft.sum()).backward()
        grads.append(leaf.grad.clone())
    assert rel_close(grads[0], grads[1], 1e-8)
