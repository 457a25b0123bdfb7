"""deftet_sparse_render (kaolin/render/mesh/deftet.py:269-417, kaolin/csrc/render/mesh/deftet_cuda.cu; SURVEY.md
§8f row 4).

CPU: the C++ oracle (+ restated torch glue) against the reference's known-answer vectors
(tests/python/kaolin/render/mesh/test_deftet.py:35-243) and against golden outputs / gradients of the
reference's pure-torch ``_naive_deftet_sparse_render`` (tests/golden/make_golden_deftet.py).
GPU: the CUDA path through the C ABI against the same vectors, against the reference's OWN CUDA kernels
(oracle/_ref/ref_C.so) and against the oracle on seeded scenes, incl. the cases the grid binning must get
right: points outside [-1,1], huge and tiny faces, more than knum hits, more hits than the shared-memory
buffer, depth ties.  face_idx: torch.equal.  Features: the reference's own tolerances (rtol = atol = 1e-4 vs
naive; 1e-5 relative vs the reference kernels).  Gradients: 1e-4 relative vs the reference kernels."""
import numpy as np
import pytest
import torch

from helpers import golden
from oracle import oracle as O

DT = {"f32": torch.float32, "f64": torch.float64}


def rel_close(a, b, tol, name=""):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (name, a.shape, b.shape)
    scale = max(float(b.abs().max()), 1e-30)
    err = float((a - b).abs().max())
    assert err <= tol * scale, f"{name}: max abs err {err:.3e} vs scale {scale:.3e}"


# ------------------------------------------------------------------ known-answer scene of the reference
def simple_scene(device, dtype):
    img = torch.tensor(
        [[[[-1., 0.], [0., -1.], [0., 1.]], [[-1., 0.], [0., 1.], [0., -1.]], [[0., -1.], [0., 1.], [1., 0.]]],
         [[[-1., -1.], [1., -1.], [-1., 1.]], [[-1., -1.], [1., -1.], [-1., 1.]], [[-1., -1.], [1., -1.], [-1., 1.]]]],
        device=device, dtype=dtype)
    z = torch.tensor([[[-2., -1., -1.], [-2.5, -3., -3.], [-2., -2., -2.]],
                      [[-2., -1., -3.], [-2., -2., -2.], [-2., -3., -1.]]], device=device, dtype=dtype)
    per_face = torch.tensor([[[[0.], [0.], [0.]], [[1.], [1.], [1.]], [[2.], [2.], [2.]]],
                             [[[3.], [3.], [3.]], [[4.], [4.], [4.]], [[5.], [5.], [5.]]]], device=device, dtype=dtype)
    per_vertex = torch.arange(18, device=device, dtype=dtype).reshape(2, 3, 3, 1)
    pix = torch.tensor(
        [[[-0.999, 0.], [-0.001, -0.998], [0.001, 0.998], [0.999, 0.], [-0.45, 0.], [0.45, 0.], [-0.999, -0.999]],
         [[-0.998, -0.999], [0.998, -0.999], [-0.999, 0.998], [-0.001, -0.], [0., -0.999], [-0.999, 0.],
          [0.001, 0.001]]], device=device, dtype=dtype)
    return img, z, [per_face, per_vertex], pix


GT_FULL_IDX = [[[0, 1, -1, -1, -1], [0, 1, -1, -1, -1], [2, -1, -1, -1, -1], [2, -1, -1, -1, -1],
                [0, 1, -1, -1, -1], [2, -1, -1, -1, -1], [-1, -1, -1, -1, -1]],
               [[0, 1, 2, -1, -1], [0, 1, 2, -1, -1], [2, 1, 0, -1, -1], [2, 1, 0, -1, -1], [0, 1, 2, -1, -1],
                [2, 1, 0, -1, -1], [-1, -1, -1, -1, -1]]]
GT_FULL_F1 = [[[0., 3., 0., 0., 0.], [1., 5., 0., 0., 0.], [7., 0., 0., 0., 0.], [8., 0., 0., 0., 0.],
               [0.825, 3.825, 0., 0., 0.], [7.175, 0., 0., 0., 0.], [0., 0., 0., 0., 0.]],
              [[9., 12., 15., 0., 0.], [10., 13., 16., 0., 0.], [17., 14., 11., 0., 0.], [16.5, 13.5, 10.5, 0., 0.],
               [9.5, 12.5, 15.5, 0., 0.], [16., 13., 10., 0., 0.], [0., 0., 0., 0., 0.]]]
GT_RESTRICTED_IDX = [[[0, -1, -1, -1, -1], [0, -1, -1, -1, -1], [2, -1, -1, -1, -1], [2, -1, -1, -1, -1],
                      [0, -1, -1, -1, -1], [2, -1, -1, -1, -1], [-1, -1, -1, -1, -1]],
                     [[0, 1, 2, -1, -1], [0, 1, -1, -1, -1], [2, 1, -1, -1, -1], [2, 1, 0, -1, -1],
                      [0, 1, -1, -1, -1], [2, 1, -1, -1, -1], [-1, -1, -1, -1, -1]]]
GT_RESTRICTED_F1 = [[[0., 0., 0., 0., 0.], [1., 0., 0., 0., 0.], [7., 0., 0., 0., 0.], [8., 0., 0., 0., 0.],
                     [0.825, 0., 0., 0., 0.], [7.175, 0., 0., 0., 0.], [0., 0., 0., 0., 0.]],
                    [[9., 12., 15., 0., 0.], [10., 13., 0., 0., 0.], [17., 14., 0., 0., 0.],
                     [16.5, 13.5, 10.5, 0., 0.], [9.5, 12.5, 0., 0., 0.], [16., 13., 0., 0., 0.],
                     [0., 0., 0., 0., 0.]]]


def check_known_answers(render, device, dtype):
    """test_deftet.py:105-243 (test_full_render, test_restricted_range); render(pix, rr, z, img, feats, knum) ->
    ((f0, f1), face_idx)."""
    img, z, feats, pix = simple_scene(device, dtype)
    for lo, gt_idx, gt_f1 in ((-4., GT_FULL_IDX, GT_FULL_F1), (-2.1, GT_RESTRICTED_IDX, GT_RESTRICTED_F1)):
        rr = torch.tensor([[[lo, 0.]]], device=device, dtype=dtype).repeat(2, 7, 1)
        (f0, f1), idx = render(pix, rr, z, img, feats, 5)
        gt_idx_t = torch.tensor(gt_idx, device=device, dtype=torch.long)
        assert torch.equal(idx, gt_idx_t)
        gt_f0 = (gt_idx_t + torch.arange(2, device=device).view(2, 1, 1) * 3).to(dtype).unsqueeze(-1)
        gt_f0[gt_idx_t == -1] = 0.
        assert torch.allclose(f0, gt_f0)
        assert torch.allclose(f1, torch.tensor(gt_f1, device=device, dtype=dtype).unsqueeze(-1), atol=3e-3, rtol=1e-5)


def oracle_render(pix, rr, z, img, feats, knum, backend=None):
    cat = torch.cat(feats, dim=-1)
    out, idx, _ = O.deftet_sparse_render_forward(pix, rr, z, img, cat, knum, 1e-8, backend=backend)
    outs, c0 = [], 0
    for f in feats:
        outs.append(out[..., c0:c0 + f.shape[-1]])
        c0 += f.shape[-1]
    return tuple(outs), idx


def golden_case(dname, rname, device="cpu"):
    g = golden("deftet_naive")
    t = lambda k: torch.from_numpy(g[f"{dname}_{k}"]).to(device)  # noqa: E731
    return dict(pix=t("pixel_coords"), z=t("face_vertices_z"), img=t("face_vertices_image"), uv=t("face_uvs"),
                rr=t(f"{rname}_render_ranges"), idx=t(f"{rname}_face_idx").long(), f_uv=t(f"{rname}_feat_uv"),
                f_mask=t(f"{rname}_feat_mask"), g_uv=t(f"{rname}_g_uv"), g_mask=t(f"{rname}_g_mask"),
                grad_image=t(f"{rname}_grad_image"), grad_uvs=t(f"{rname}_grad_uvs"),
                grad_mask=t(f"{rname}_grad_mask"))


KNUM_GOLDEN = 24


@pytest.mark.parametrize("dname", ["f32", "f64"])
def test_oracle_known_answers(dname):
    check_known_answers(oracle_render, "cpu", DT[dname])


@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("rname", ["full", "center"])
def test_oracle_matches_naive_golden(dname, rname):
    c = golden_case(dname, rname)
    mask = torch.ones_like(c["uv"][..., :1])
    feats = torch.cat([c["uv"], mask], -1).contiguous()
    out, idx, w = O.deftet_sparse_render_forward(c["pix"], c["rr"], c["z"], c["img"], feats, KNUM_GOLDEN)
    assert torch.equal(idx, c["idx"])
    assert torch.allclose(out[..., :2], c["f_uv"], rtol=1e-4, atol=1e-4)          # test_deftet.py:429-431
    assert torch.allclose(out[..., 2:], c["f_mask"], rtol=1e-4, atol=1e-4)
    gi, gf = O.deftet_sparse_render_backward(torch.cat([c["g_uv"], c["g_mask"]], -1).contiguous(), idx, w, c["img"],
                                             feats)
    assert torch.allclose(gi, c["grad_image"], rtol=5e-3, atol=5e-3)              # test_deftet.py:490-495
    assert torch.allclose(gf[..., :2], c["grad_uvs"], rtol=1e-3, atol=1e-3)
    assert torch.allclose(gf[..., 2:], c["grad_mask"], rtol=1e-3, atol=1e-3)


def test_oracle_first_knum_in_face_order():
    """More hits than knum: the first knum faces in mesh order survive (deftet.py:112-116), then depth-sorted."""
    dtype = torch.float64
    tri = torch.tensor([[-1., -1.], [1., -1.], [-1., 1.]], dtype=dtype)
    img = tri.reshape(1, 1, 3, 2).repeat(1, 6, 1, 1)
    z = torch.tensor([-1., -5., -2., -4., -3., -6.], dtype=dtype).reshape(1, 6, 1).repeat(1, 1, 3)
    pix = torch.tensor([[[-0.5, -0.5]]], dtype=dtype)
    rr = torch.tensor([[[-10., 0.]]], dtype=dtype)
    feats = torch.arange(6, dtype=dtype).reshape(1, 6, 1, 1).repeat(1, 1, 3, 1)
    out, idx, _ = O.deftet_sparse_render_forward(pix, rr, z, img, feats, 4)
    assert idx.tolist() == [[[0, 2, 3, 1]]]               # faces 0..3 kept, ordered by depth -1 > -2 > -4 > -5
    assert out.reshape(-1).tolist() == [0., 2., 3., 1.]


# ----------------------------------------------------------------------------------------- GPU
def cuda_render(pix, rr, z, img, feats, knum):
    from kaolin_b200.render.mesh import deftet_sparse_render
    return deftet_sparse_render(pix, rr, z, img, feats, knum)


@pytest.mark.gpu
@pytest.mark.parametrize("dname", ["f32", "f64"])
def test_cuda_known_answers(dname):
    check_known_answers(cuda_render, "cuda", DT[dname])
    # a single (concatenated) tensor returns a tensor, as in the reference (cat_features=True)
    img, z, feats, pix = simple_scene("cuda", DT[dname])
    rr = torch.tensor([[[-4., 0.]]], device="cuda", dtype=DT[dname]).repeat(2, 7, 1)
    out, idx = cuda_render(pix, rr, z, img, torch.cat(feats, -1), 5)
    assert isinstance(out, torch.Tensor) and out.shape == (2, 7, 5, 2)
    assert torch.equal(idx, torch.tensor(GT_FULL_IDX, device="cuda"))


@pytest.mark.gpu
@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("rname", ["full", "center"])
def test_cuda_matches_naive_golden(dname, rname):
    c = golden_case(dname, rname, "cuda")
    img = c["img"].clone().requires_grad_(True)
    uv = c["uv"].clone().requires_grad_(True)
    mask = torch.ones_like(uv[..., :1]).requires_grad_(True)
    (f_uv, f_mask), idx = cuda_render(c["pix"], c["rr"], c["z"], img, [uv, mask], KNUM_GOLDEN)
    assert torch.equal(idx, c["idx"])
    assert torch.allclose(f_uv, c["f_uv"], rtol=1e-4, atol=1e-4)
    assert torch.allclose(f_mask, c["f_mask"], rtol=1e-4, atol=1e-4)
    torch.autograd.backward([f_uv, f_mask], [c["g_uv"], c["g_mask"]])
    assert torch.allclose(img.grad, c["grad_image"], rtol=5e-3, atol=5e-3)
    assert torch.allclose(uv.grad, c["grad_uvs"], rtol=1e-3, atol=1e-3)
    assert torch.allclose(mask.grad, c["grad_mask"], rtol=1e-3, atol=1e-3)


def random_scene(B, F, P, scale, seed, dtype, spread=1.0, device="cuda"):
    g = torch.Generator().manual_seed(seed)
    centre = (torch.rand(B, F, 1, 2, generator=g, dtype=torch.float64) * 2 - 1) * spread
    img = centre + (torch.rand(B, F, 3, 2, generator=g, dtype=torch.float64) - 0.5) * scale
    z = -(torch.rand(B, F, 3, generator=g, dtype=torch.float64) * 4 + 0.5)
    pix = (torch.rand(B, P, 2, generator=g, dtype=torch.float64) * 2 - 1) * spread * 1.1
    rr = torch.stack([-(torch.rand(B, P, generator=g, dtype=torch.float64) * 3 + 1.5),
                      torch.zeros(B, P, dtype=torch.float64)], -1)
    feats = torch.rand(B, F, 3, 4, generator=g, dtype=torch.float64)
    return [t.to(dtype).to(device).contiguous() for t in (img, z, pix, rr, feats)]


def check_vs_backend(backend, B, F, P, scale, knum, seed, dtype, spread=1.0, grads=True):
    """CUDA path vs `backend` (None = C++ oracle on CPU, else ref_C on the GPU) on the same tensors."""
    import kaolin_b200 as kal
    img, z, pix, rr, feats = random_scene(B, F, P, scale, seed, dtype, spread)
    mv = (lambda t: t.cpu()) if backend is None else (lambda t: t)
    r_out, r_idx, r_w = O.deftet_sparse_render_forward(mv(pix), mv(rr), mv(z), mv(img), mv(feats), knum, 1e-8,
                                                       backend=backend)
    il, fl = img.clone().requires_grad_(True), feats.clone().requires_grad_(True)
    out, idx = kal.render.mesh.deftet_sparse_render(pix, rr, z, il, fl, knum)
    assert torch.equal(idx.cpu(), r_idx.cpu()), f"face_idx differs on {(idx.cpu() != r_idx.cpu()).sum().item()} slots"
    rel_close(out, r_out, 1e-5 if dtype == torch.float32 else 1e-12, "features")
    # native-op shape (face order, depth / w0 / w1 with the reference's fills)
    bb = torch.cat([img.min(dim=2)[0], img.max(dim=2)[0]], dim=2).contiguous()
    n_idx, n_depth, n_w0, n_w1 = kal._C.render.mesh.deftet_sparse_render_forward_cuda(z, img, bb, pix, rr, knum, 1e-8)
    mesh = O.OracleC.render.mesh if backend is None else backend.render.mesh
    q_idx, q_depth, q_w0, q_w1 = mesh.deftet_sparse_render_forward_cuda(mv(z), mv(img), mv(bb), mv(pix), mv(rr),
                                                                        knum, 1e-8)
    assert torch.equal(n_idx.cpu(), q_idx.cpu())
    assert torch.equal(n_depth.cpu(), q_depth.cpu()) and torch.equal(n_w0.cpu(), q_w0.cpu())
    assert torch.equal(n_w1.cpu(), q_w1.cpu())
    if grads:
        g = torch.Generator().manual_seed(seed + 1)
        go = torch.rand(out.shape, generator=g, dtype=torch.float64).to(dtype).cuda()
        out.backward(go)
        r_gi, r_gf = O.deftet_sparse_render_backward(mv(go), r_idx, r_w, mv(img), mv(feats), 1e-8, backend=backend)
        gtol = 1e-4 if dtype == torch.float32 else 1e-9
        rel_close(fl.grad, r_gf, gtol, "grad_face_features")
        rel_close(il.grad, r_gi, gtol, "grad_face_vertices_image")
    return idx


@pytest.fixture(scope="module")
def ref_C():
    m = O.load_ref_C()
    if m is None or not hasattr(m.render.mesh, "deftet_sparse_render_forward_cuda"):
        pytest.skip("oracle/_ref/ref_C.so (with deftet) not built")
    return m


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("B,F,P,scale,knum", [(1, 1, 1, 2.0, 1), (2, 300, 257, 0.3, 8), (3, 5000, 1025, 0.08, 30),
                                              (1, 20000, 4096, 0.05, 20), (2, 900, 400, 1.5, 300)])
def test_cuda_vs_reference_cuda_kernels(ref_C, dtype, B, F, P, scale, knum):
    check_vs_backend(ref_C, B, F, P, scale, knum, 11 + F, dtype)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_cuda_vs_oracle_bit_exact_selection(dtype):
    check_vs_backend(None, 2, 1500, 300, 0.2, 12, 5, dtype)


@pytest.mark.gpu
def test_cuda_points_and_faces_outside_the_grid(ref_C):
    """Geometry spread over [-3,3]^2: everything beyond [-1,1] lands in border cells of the grid."""
    check_vs_backend(ref_C, 2, 3000, 2000, 0.4, 16, 77, torch.float32, spread=3.0)


@pytest.mark.gpu
def test_cuda_more_hits_than_knum_and_than_the_buffer(ref_C):
    """2500 screen-filling faces, up to ~360 hits per point: knum 7 (64-entry buffer -> repeated compaction),
    knum 200 (256-entry buffer: overflows once for the deepest points) and knum 1000 (nothing truncated)."""
    for knum in (7, 200, 1000):
        idx = check_vs_backend(ref_C, 1, 2500, 64, 3.0, knum, 3, torch.float32, grads=(knum == 7))
        assert int((idx >= 0).sum(-1).max()) == min(knum, 359)


@pytest.mark.gpu
def test_cuda_depth_ties_keep_face_order():
    import kaolin_b200 as kal
    tri = torch.tensor([[-1., -1.], [1., -1.], [-1., 1.]], device="cuda")
    img = tri.reshape(1, 1, 3, 2).repeat(1, 5, 1, 1).contiguous()
    z = torch.tensor([-2., -1., -2., -1., -2.], device="cuda").reshape(1, 5, 1).repeat(1, 1, 3).contiguous()
    feats = torch.arange(5, device="cuda", dtype=torch.float32).reshape(1, 5, 1, 1).repeat(1, 1, 3, 1).contiguous()
    pix = torch.tensor([[[-0.5, -0.5]]], device="cuda")
    rr = torch.tensor([[[-10., 0.]]], device="cuda")
    out, idx = kal.render.mesh.deftet_sparse_render(pix, rr, z, img, feats, 6)
    assert idx.tolist() == [[[1, 3, 0, 2, 4, -1]]]
    assert out.reshape(-1).tolist() == [1., 3., 0., 2., 4., 0.]


@pytest.mark.gpu
def test_cuda_empty_inputs_and_errors():
    import kaolin_b200 as kal
    img, z, pix, rr, feats = random_scene(2, 50, 9, 0.5, 1, torch.float32)
    out, idx = kal.render.mesh.deftet_sparse_render(pix[:, :0], rr[:, :0], z, img, feats, 4)
    assert out.shape == (2, 0, 4, 4) and idx.shape == (2, 0, 4)
    out, idx = kal.render.mesh.deftet_sparse_render(pix, rr, z[:, :0], img[:, :0], feats[:, :0], 4)
    assert bool((idx == -1).all()) and bool((out == 0).all())
    nan_pix = pix.clone()
    nan_pix[0, 0, 0] = float("nan")
    out, idx = kal.render.mesh.deftet_sparse_render(nan_pix, rr, z, img, feats, 4)
    assert bool((idx[0, 0] == -1).all())
    with pytest.raises(RuntimeError):
        kal.render.mesh.deftet_sparse_render(pix.cpu(), rr.cpu(), z.cpu(), img.cpu(), feats.cpu(), 4)
    with pytest.raises(RuntimeError):
        kal.render.mesh.deftet_sparse_render(pix, rr[:, :3], z, img, feats, 4)
    with pytest.raises(RuntimeError):
        kal.render.mesh.deftet_sparse_render(pix, rr, z, img, feats, 5000)          # knum above the library limit
    with pytest.raises(RuntimeError):
        kal._C.render.mesh.deftet_sparse_render_forward_cuda(z, img, img[:, :, 0], pix, rr, 4, 1e-8)
    torch.cuda.synchronize()


@pytest.mark.gpu
def test_cuda_gradients_only_for_requested_inputs():
    import kaolin_b200 as kal
    img, z, pix, rr, feats = random_scene(1, 400, 200, 0.4, 9, torch.float32)
    pl, rl, zl = (t.clone().requires_grad_(True) for t in (pix, rr, z))
    fl = feats.clone().requires_grad_(True)
    out, idx = kal.render.mesh.deftet_sparse_render(pl, rl, zl, img, fl, 10)
    out.sum().backward()
    # test_deftet.py:487-489: no gradient w.r.t. pixel_coords, render_ranges, face_vertices_z
    assert pl.grad is None and rl.grad is None and zl.grad is None
    assert fl.grad is not None and float(fl.grad.abs().sum()) > 0
    assert not idx.requires_grad


@pytest.mark.gpu
def test_cuda_backward_with_only_non_differentiable_inputs_requiring_grad():
    """Only face_vertices_z (or pixel_coords / render_ranges) requires grad: the outputs require grad, backward is
    called, and — like the reference, deftet.py:337 — returns None for those inputs without error."""
    import kaolin_b200 as kal
    img, z, pix, rr, feats = random_scene(1, 300, 100, 0.4, 4, torch.float32)
    for which in range(3):
        leaves = [t.clone().requires_grad_(i == which) for i, t in enumerate((pix, rr, z))]
        out, idx = kal.render.mesh.deftet_sparse_render(leaves[0], leaves[1], leaves[2], img, feats, 6)
        assert out.requires_grad
        out.sum().backward()
        assert all(t.grad is None for t in leaves)
