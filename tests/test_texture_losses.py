"""texture_mapping (kaolin/render/mesh/utils.py:23-76) and mask_iou
(kaolin/metrics/render.py:18-40; row 3).

CPU: the oracle restatements against the reference's known-answer vectors
(tests/python/kaolin/render/mesh/test_utils.py:26-112, tests/python/kaolin/metrics/test_render.py:24-53).
GPU: the CUDA kernels (through the C ABI) against the same vectors (torch.equal, as the reference's tests assert)
and against the oracle on seeded inputs, forward and backward.  Tolerances: 1e-5 relative forward, 1e-4 gradients
(fp32); 1e-12 / 1e-10 (fp64)."""
import pytest
import torch

from oracle import oracle as O

DT = {"f32": torch.float32, "f64": torch.float64}


def rel_close(a, b, tol):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    scale = max(float(b.abs().max()), 1e-30)
    return a.shape == b.shape and float((a - b).abs().max()) <= tol * scale


# ------------------------------------------------------------------ known-answer vectors of the reference
def _texture_map_1d(device, dtype):
    l1 = torch.tensor([[11.0, 12.0, 13.0, 14.0, 15.0], [21.0, 22.0, 23.0, 24.0, 25.0],
                       [31.0, 32.0, 33.0, 34.0, 35.0], [41.0, 42.0, 43.0, 44.0, 45.0]], device=device, dtype=dtype)
    return torch.stack((l1, l1 + 100)).unsqueeze(1)


def _texture_map_3d(device, dtype):
    t = _texture_map_1d(device, dtype)
    return torch.cat((t, -t, t), dim=1)


def _sparse_coords(device, dtype):
    c = torch.tensor([[0.0, 0.0], [1.0, 1.0], [0, 1.0], [0.5, 0.5]], device=device, dtype=dtype)
    return torch.stack((c, torch.flip(c, dims=(0,))))


def _dense_coords(device, dtype):
    c = torch.tensor([[[0.0, 0.0], [0.25, 0.0], [0.5, 0.0], [0.75, 0.0], [1.0, 0.0]],
                      [[0.0, 1 / 8], [0.25, 1 / 8], [0.5, 1 / 8], [0.75, 1 / 8], [1.0, 1 / 8]]],
                     device=device, dtype=dtype)
    return torch.stack((c, torch.flip(c, dims=(0,))))


def _known_answers(fn, device, dtype):
    """The three tests of TestTextureMapping, for both modes."""
    for mode in ("nearest", "bilinear"):
        base = [[41, 15, 11, 33], [133, 111, 115, 141]] if mode == "nearest" else \
            [[41, 15, 11, 28], [128, 111, 115, 141]]
        e1 = torch.tensor(base, device=device).to(dtype)
        out = fn(_sparse_coords(device, dtype), _texture_map_1d(device, dtype), mode=mode)
        assert out.shape == (2, 4, 1) and out.dtype == dtype
        assert torch.equal(out, e1.unsqueeze(-1))
        out = fn(_sparse_coords(device, dtype), _texture_map_3d(device, dtype), mode=mode)
        assert out.shape == (2, 4, 3)
        assert torch.equal(out, torch.stack([e1, -e1, e1], dim=-1))
        eb = torch.tensor([41., 42., 43., 44., 45.] if mode == "nearest" else [41., 41.75, 43., 44.25, 45.],
                          device=device, dtype=dtype)
        e = torch.stack([eb, eb + 100], dim=0)
        e = torch.stack([e, -e, e], dim=-1).reshape(2, 1, -1, 3).repeat(1, 2, 1, 1)
        out = fn(_dense_coords(device, dtype), _texture_map_3d(device, dtype), mode=mode)
        assert torch.equal(out, e)


def _iou_masks(device, dtype):
    lhs = torch.tensor([[[0., 0.2, 0.1, 1.], [0.5, 0.5, 0.9, 0.9], [0., 1., 1., 0.9], [0.8, 0.7, 0.2, 0.1]],
                        [[1., 1., 1., 1.], [1., 1., 1., 1.], [1., 1., 1., 1.], [1., 1., 1., 1.]]],
                       dtype=dtype, device=device)
    rhs = torch.tensor([[[0.1, 0.3, 0.3, 0.9], [0.5, 0.5, 1., 0.3], [0., 0.9, 0.9, 0.8], [1., 1., 0., 0.]],
                        [[0.3, 0.6, 0.7, 0.7], [0.8, 0.9, 0.9, 1.], [1., 0.9, 0.9, 0.5], [0.8, 0.7, 0.8, 0.5]]],
                       dtype=dtype, device=device)
    return lhs, rhs


@pytest.mark.parametrize("dname", ["f32", "f64"])
def test_oracle_texture_mapping_known_answers(dname):
    _known_answers(O.texture_mapping, "cpu", DT[dname])


@pytest.mark.parametrize("dname", ["f32", "f64"])
def test_oracle_mask_iou_known_answer(dname):
    lhs, rhs = _iou_masks("cpu", DT[dname])
    loss = O.mask_iou(lhs, rhs)
    assert torch.allclose(loss, torch.tensor([0.3105], dtype=DT[dname]))      # test_render.py:50-53


# ----------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("dname", ["f32", "f64"])
def test_cuda_texture_mapping_known_answers(dname):
    from kaolin_b200.render.mesh import texture_mapping
    _known_answers(texture_mapping, "cuda", DT[dname])


@pytest.mark.gpu
@pytest.mark.parametrize("dname", ["f32", "f64"])
def test_cuda_mask_iou_known_answer(dname):
    from kaolin_b200.metrics.render import mask_iou
    lhs, rhs = _iou_masks("cuda", DT[dname])
    loss = mask_iou(lhs, rhs)
    assert loss.dim() == 0
    assert torch.allclose(loss, torch.tensor([0.3105], dtype=DT[dname], device="cuda"))
    assert rel_close(loss, O.mask_iou(lhs, rhs), 1e-6 if dname == "f32" else 1e-14)


def _uv_scene(dtype, B, H, W, C, Ht, Wt, seed, shared):
    g = torch.Generator().manual_seed(seed)
    uv = torch.rand(B, H, W, 2, generator=g, dtype=torch.float64) * 1.3 - 0.15     # some outside [0,1]
    uv[0, 0, 0] = torch.tensor([0.0, 1.0])
    uv[0, 0, 1] = torch.tensor([1.0, 0.0])
    tex = torch.rand(1 if shared else B, C, Ht, Wt, generator=g, dtype=torch.float64) * 1.4 - 0.2
    mask = (torch.rand(B, H, W, 1, generator=g, dtype=torch.float64) * 1.2)
    gout = torch.randn(B, H, W, C, generator=g, dtype=torch.float64)
    return [t.to(dtype).cuda() for t in (uv, tex, mask, gout)]


@pytest.mark.gpu
@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("mode", ["nearest", "bilinear"])
@pytest.mark.parametrize("shared", [False, True])
@pytest.mark.parametrize("shade", [False, True])
def test_cuda_texture_mapping_matches_oracle(dname, mode, shared, shade):
    """Dense (B,H,W,2) coordinates incl. out-of-range values, per-view or shared texture, with and without the
    clamp(image * mask, 0, 1) epilogue; gradients w.r.t. texture, coordinates and mask."""
    from kaolin_b200.render.mesh import texture_mapping
    dtype = DT[dname]
    B, H, W, C, Ht, Wt = 3, 37, 53, 3, 29, 41
    uv, tex, mask, gout = _uv_scene(dtype, B, H, W, C, Ht, Wt, 5, shared)
    # reference in fp64 on the same (dtype-rounded) inputs
    leaves = [t.clone().requires_grad_(True) for t in (uv, tex, mask)]
    refs = [t.double().clone().requires_grad_(True) for t in (uv, tex, mask)]
    got = texture_mapping(leaves[0], leaves[1], mode=mode, mask=leaves[2] if shade else None)
    want = O.texture_mapping(refs[0], refs[1].expand(B, -1, -1, -1), mode=mode)
    if shade:
        want = torch.clamp(want * refs[2], 0., 1.)
    assert got.dtype == dtype and got.shape == (B, H, W, C)
    ftol, gtol = (1e-5, 1e-4) if dname == "f32" else (1e-12, 1e-10)
    if mode == "nearest" and dname == "f32":
        # a coordinate that rounds to xx.5 in fp32 may pick the neighbouring texel in fp64: compare where both
        # agree on the texel, and require that to be nearly everywhere
        same = ((got.detach().double() - want.detach()).abs().amax(-1) <= ftol * float(want.abs().max()))
        assert same.double().mean() > 0.999
    else:
        assert rel_close(got, want, ftol)
    got.backward(gout)
    want.backward(gout.double())
    if not (mode == "nearest" and dname == "f32"):
        assert rel_close(leaves[1].grad, refs[1].grad, gtol)
        if mode == "bilinear":
            assert rel_close(leaves[0].grad, refs[0].grad, gtol * 10 if dname == "f32" else gtol)
        else:
            assert float(leaves[0].grad.abs().max()) == 0.0
        if shade:
            assert rel_close(leaves[2].grad, refs[2].grad, gtol)


@pytest.mark.gpu
def test_cuda_texture_mapping_sparse_empty_and_errors():
    from kaolin_b200.render.mesh import texture_mapping
    tex = torch.rand(2, 3, 8, 8, device="cuda")
    pts = torch.rand(2, 17, 2, device="cuda")
    out = texture_mapping(pts, tex, mode="bilinear")
    assert out.shape == (2, 17, 3) and rel_close(out, O.texture_mapping(pts.cpu(), tex.cpu(), "bilinear"), 1e-5)
    assert texture_mapping(pts[:, :0], tex).shape == (2, 0, 3)
    with pytest.raises(ValueError):
        texture_mapping(pts, tex, mode="cubic")
    with pytest.raises(RuntimeError):
        texture_mapping(pts.cpu(), tex.cpu())
    with pytest.raises(RuntimeError):
        texture_mapping(pts, tex[:, :, 0])
    with pytest.raises(RuntimeError):
        texture_mapping(pts, torch.rand(3, 3, 8, 8, device="cuda"))
    with pytest.raises(RuntimeError):
        texture_mapping(pts.half(), tex.half())


@pytest.mark.gpu
@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("shape", [(1, 1, 1), (3, 64, 64), (2, 91, 131), (5, 256, 256)])
def test_cuda_mask_iou_matches_oracle(dname, shape):
    from kaolin_b200.metrics.render import mask_iou
    dtype = DT[dname]
    g = torch.Generator().manual_seed(shape[1])
    lhs = torch.rand(shape, generator=g, dtype=torch.float64).to(dtype).cuda()
    rhs = (torch.rand(shape, generator=g, dtype=torch.float64) > 0.4).to(dtype).cuda()
    a, b = lhs.clone().requires_grad_(True), rhs.clone().requires_grad_(True)
    c, d = lhs.double().clone().requires_grad_(True), rhs.double().clone().requires_grad_(True)
    got, want = mask_iou(a, b), O.mask_iou(c, d)
    ftol, gtol = (1e-5, 1e-4) if dname == "f32" else (1e-12, 1e-10)
    assert rel_close(got, want, ftol)
    (got * 3.0).backward()
    (want * 3.0).backward()
    assert rel_close(a.grad, c.grad, gtol) and rel_close(b.grad, d.grad, gtol)
    # deterministic: a second evaluation is bit-identical (two-stage reduction, no float atomics)
    assert torch.equal(mask_iou(lhs, rhs), got.detach())


@pytest.mark.gpu
def test_cuda_mask_iou_edge_cases():
    from kaolin_b200.metrics.render import mask_iou
    z = torch.zeros(2, 8, 8, device="cuda")
    assert float(mask_iou(z, z)) == 1.0                      # 0 / (0 + 1e-10)
    with pytest.raises(AssertionError):
        mask_iou(z, z[:1])
    with pytest.raises(RuntimeError):
        mask_iou(z.cpu(), z.cpu())
    only_lhs = torch.rand(2, 8, 8, device="cuda", requires_grad=True)
    mask_iou(only_lhs, z + 1).backward()
    assert only_lhs.grad is not None and torch.isfinite(only_lhs.grad).all()
