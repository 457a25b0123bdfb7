"""``dibr_rasterization`` as ONE native call (both list sets from one read, coverage maps handed from the rasterizer's
epilogue to the soft-mask kernel) against the reference's composition ``rasterize(valid_faces = normals_z >= 0)`` +
``dibr_soft_mask`` (kaolin/render/mesh/dibr.py:119-209) on inputs that stress the hand-over: ragged image sizes (partial
tiles), NaN depths (tiles that fall back to the in-order scan), degenerate / off-screen / screen-filling faces, culled
faces, fp64 (no hand-over), lists of feature tensors.  Forward values must be identical (test_dibr.py:527-529 asserts the
same of the reference), gradients within the tolerance `north_star` states."""
import pytest
import torch

from helpers import norm_close

pytestmark = pytest.mark.gpu


def _scene(B, F, seed, dtype, nan_faces=0, big=False):
    g = torch.Generator().manual_seed(seed)
    ctr = torch.rand(B, F, 1, 2, generator=g, dtype=torch.float64) * 2.4 - 1.2        # some faces off screen
    size = 0.6 if big else 0.08
    img = ctr + (torch.rand(B, F, 3, 2, generator=g, dtype=torch.float64) - 0.5) * size
    img[:, ::17, 2] = img[:, ::17, 1]                                                   # degenerate (two equal vertices)
    img[:, 5::29] = img[:, 5::29, :1]                                                   # a point
    z = -1.0 - 2.0 * torch.rand(B, F, 3, generator=g, dtype=torch.float64)
    if nan_faces:
        z[:, :nan_faces, 0] = float("nan")
    nz = torch.rand(B, F, generator=g, dtype=torch.float64) - 0.3                        # ~30 % culled
    feats = [torch.rand(B, F, 3, c, generator=g, dtype=torch.float64) for c in (3, 2, 4)]
    to = lambda t: t.to(dtype).cuda()                                                    # noqa: E731
    return to(z), to(img), [to(f) for f in feats], to(nz)


@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("H,W,B,F,nan_faces,big", [
    (35, 31, 2, 300, 0, False), (1, 1, 1, 40, 0, True), (16, 240, 3, 500, 0, False), (67, 130, 2, 800, 6, False),
    (97, 64, 1, 60, 3, True), (128, 128, 2, 3000, 0, False)])
def test_fused_equals_rasterize_plus_soft_mask(dname, H, W, B, F, nan_faces, big):
    import kaolin_b200 as kal
    dtype = torch.float32 if dname == "f32" else torch.float64
    z, img0, feats0, nz = _scene(B, F, 1000 + H * W + F, dtype, nan_faces, big)
    g = torch.Generator().manual_seed(7)
    g_out = [torch.rand(B, H, W, f.shape[-1], generator=g, dtype=torch.float64).to(dtype).cuda() for f in feats0]
    g_soft = torch.rand(B, H, W, generator=g, dtype=torch.float64).to(dtype).cuda()
    res = []
    for fused in (False, True):
        img = img0.clone().requires_grad_(True)
        feats = [f.clone().requires_grad_(True) for f in feats0]
        if fused:
            out, soft, idx = kal.render.mesh.dibr_rasterization(H, W, z, img, feats, nz, knum=12)
        else:
            out, idx = kal.render.mesh.rasterize(H, W, z, img, feats, nz >= 0.)
            soft = kal.render.mesh.dibr_soft_mask(img, idx, knum=12)
        torch.autograd.backward(list(out) + [soft], g_out + [g_soft])
        res.append((out, soft, idx, img.grad, [f.grad for f in feats]))
    (o0, s0, i0, gi0, gf0), (o1, s1, i1, gi1, gf1) = res
    assert torch.equal(i0, i1)
    for a, b in zip(o0, o1):
        assert torch.equal(a, b) or torch.equal(torch.nan_to_num(a, nan=-7.0), torch.nan_to_num(b, nan=-7.0))
    assert torch.equal(s0, s1)
    finite = torch.isfinite(gi0) & torch.isfinite(gi1)
    assert torch.equal(torch.isfinite(gi0), torch.isfinite(gi1))
    norm_close(torch.where(finite, gi1, torch.zeros_like(gi1)), torch.where(finite, gi0, torch.zeros_like(gi0)), 1e-4,
               "grad face_vertices_image")
    for a, b in zip(gf0, gf1):
        fin = torch.isfinite(a) & torch.isfinite(b)
        norm_close(torch.where(fin, b, torch.zeros_like(b)), torch.where(fin, a, torch.zeros_like(a)), 1e-4,
                   "grad face_features")
