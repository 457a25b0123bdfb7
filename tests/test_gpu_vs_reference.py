"""GPU parity against (1) the reference's own CUDA kernels (ref_C) and (2) the C++ oracle, on synthetic
workloads of BASELINE.json `configs`, random soups and edge cases.  All calls go through the C-ABI.

Tolerances (BASELINE.json north_star): face_idx bit-exact; features 1e-5 relative; soft mask and gradients
1e-4 relative."""
import math

import pytest
import torch

from helpers import elem_close

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def kal():
    import kaolin_b200
    return kaolin_b200


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    return oracle


@pytest.fixture(scope="module")
def ref_C(O):
    m = O.load_ref_C()
    if m is None:
        pytest.skip("oracle/_ref/ref_C.so not built")
    return m


def rel_close(a, b, rel, name=""):
    a, b = a.double(), b.double()
    if b.numel() == 0:
        assert a.numel() == 0
        return
    scale = max(1.0, float(b.abs().max()))
    err = float((a - b).abs().max())
    assert err <= rel * scale, f"{name}: max abs err {err:.3e} > {rel} * {scale:.3e}"


def to_dev(d, dtype=torch.float32):
    out = dict(d)
    out["face_vertices_z"] = d["face_vertices_z"].to("cuda", dtype)
    out["face_vertices_image"] = d["face_vertices_image"].to("cuda", dtype)
    out["face_features"] = [f.to("cuda", dtype) for f in d["face_features"]]
    out["face_normals_z"] = None if d["face_normals_z"] is None else d["face_normals_z"].to("cuda", dtype)
    return out


def full_pipeline_check(kal, O, backend, d, knum=30, boxlen=0.02, sigmainv=7000., multiplier=1000., eps=1e-8,
                        check_side_buffers=True, grad_rel=1e-4):
    """Forward + backward of rasterize and dibr_soft_mask vs `backend` (ref_C on GPU tensors, or the CPU
    oracle when backend is None) on identical inputs."""
    Hh, Ww = d["H"], d["W"]
    z, img = d["face_vertices_z"], d["face_vertices_image"]
    feats = torch.cat(d["face_features"], -1).contiguous()
    valid = None if d["face_normals_z"] is None else (d["face_normals_z"] >= 0.)
    img_l = img.clone().requires_grad_(True)
    feats_l = feats.clone().requires_grad_(True)
    out_feat, face_idx = kal.render.mesh.rasterize(Hh, Ww, z, img_l, feats_l, valid, multiplier, eps)
    soft = kal.render.mesh.dibr_soft_mask(img_l, face_idx, sigmainv, boxlen, knum, multiplier)
    g = torch.Generator(device="cpu").manual_seed(77)
    g_feat = torch.rand(out_feat.shape, generator=g, dtype=torch.float64).to(out_feat)
    g_soft = torch.rand(soft.shape, generator=g, dtype=torch.float64).to(soft)
    # ---- reference
    cpu = backend is None
    mv = (lambda t: t.detach().cpu()) if cpu else (lambda t: t.detach())
    r_feat, r_idx, r_w = O.rasterize_forward(Hh, Ww, mv(z), mv(img), mv(feats), None if valid is None else mv(valid),
                                             multiplier, eps, backend=backend)
    r_soft, r_img_m, r_prob, r_cidx, r_ctype = O.dibr_soft_mask_forward(mv(img), r_idx, sigmainv, boxlen, knum,
                                                                         multiplier, backend=backend)
    assert torch.equal(face_idx.cpu(), r_idx.cpu()), \
        f"face_idx mismatch on {(face_idx.cpu() != r_idx.cpu()).sum().item()} pixels"
    # deterministic per-pixel outputs: ELEMENTWISE relative (|a-b| <= rel*|b| + tiny), worst pixel reported
    elem_close(out_feat, r_feat, 1e-5, "features", atol=1e-6 if out_feat.dtype == torch.float32 else 1e-12)
    elem_close(soft, r_soft, 1e-4, "soft_mask", atol=1e-6 if soft.dtype == torch.float32 else 1e-12)
    # ---- backward: raster and soft-mask gradients separately, then through autograd together
    r_gi, r_gf = O.rasterize_backward(mv(g_feat), r_idx, r_w, mv(img), mv(feats), eps, backend=backend,
                                      interpolated_features=r_feat)
    r_gs = O.dibr_soft_mask_backward(mv(g_soft), r_soft, r_img_m, r_idx, r_prob, r_cidx, r_ctype, sigmainv,
                                     multiplier, backend=backend)
    (gi_r, gf_r) = torch.autograd.grad(out_feat, (img_l, feats_l), g_feat, retain_graph=True)
    (gi_s,) = torch.autograd.grad(soft, (img_l,), g_soft, retain_graph=True)
    rel_close(gf_r.cpu(), r_gf.cpu(), grad_rel, "grad_face_features")
    rel_close(gi_r.cpu(), r_gi.cpu(), grad_rel, "grad_face_vertices_image (raster)")
    rel_close(gi_s.cpu(), r_gs.cpu(), grad_rel, "grad_face_vertices_image (soft mask)")
    if check_side_buffers:
        img_m = (img * multiplier).contiguous()
        pmin, pmax = img_m.min(dim=-2)[0], img_m.max(dim=-2)[0]
        bb = torch.cat([pmin - boxlen * multiplier, pmax + boxlen * multiplier], -1).contiguous()
        s2, p2, c2, t2 = kal._C.render.mesh.dibr_soft_mask_forward_cuda(img_m, bb, face_idx, sigmainv, knum, multiplier)
        # reference-shaped vs fused entry point: exact divisions / double product vs reciprocal form / fp32 product
        assert torch.allclose(s2, soft, rtol=1e-5, atol=1e-6) if soft.dtype == torch.float32 else torch.equal(s2, soft)
        assert torch.equal(c2.cpu(), r_cidx.cpu()), "close_face_idx mismatch"
        rel_close(p2.cpu(), r_prob.cpu(), 1e-5, "close_face_prob")
        mism = (t2.cpu() != r_ctype.cpu()).float().mean().item()
        assert mism <= (0.0 if not cpu else 0.01), f"dist_type mismatch fraction {mism}"
    return face_idx, soft


# ----------------------------------------------------------------------------- vs ref_C (same GPU)
def test_c1_vs_ref_C(kal, O, ref_C):
    """config 1: icosphere 1280 faces, B=1, 256x256, C=3, culling."""
    from kaolin_b200 import synth
    full_pipeline_check(kal, O, ref_C, to_dev(synth.make_inputs("c1")))


def test_c2_vs_ref_C(kal, O, ref_C):
    """config 2 at full size: 20k-face soups, B=16, 512x512, C=9, no culling."""
    from kaolin_b200 import synth
    full_pipeline_check(kal, O, ref_C, to_dev(synth.make_inputs("c2")))


def test_c3_two_views_vs_ref_C(kal, O, ref_C):
    """config 3 shape (100k faces, 1024x1024) on 2 of the 64 views — the brute-force reference needs
    ~1e11 pixel-face tests per view."""
    from kaolin_b200 import synth
    full_pipeline_check(kal, O, ref_C, to_dev(synth.make_inputs("c3", batch=2, views=[0, 37])))


def test_c1_double_vs_ref_C(kal, O, ref_C):
    from kaolin_b200 import synth
    full_pipeline_check(kal, O, ref_C, to_dev(synth.make_inputs("c1"), torch.float64), grad_rel=1e-9)


def random_soup(B, F, scale, seed, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    ctr = torch.rand(B, F, 1, 2, generator=g) * 2.2 - 1.1          # some faces off-screen
    img = ctr + (torch.rand(B, F, 3, 2, generator=g) - 0.5) * scale
    z = -1.0 - 2 * torch.rand(B, F, 3, generator=g)
    feat = torch.rand(B, F, 3, 4, generator=g)
    nz = torch.rand(B, F, generator=g) - 0.3
    return dict(face_vertices_z=z.to("cuda", dtype), face_vertices_image=img.to("cuda", dtype),
                face_features=[feat.to("cuda", dtype)], face_normals_z=nz.to("cuda", dtype))


@pytest.mark.parametrize("Hh,Ww", [(64, 64), (37, 53), (16, 240), (1, 1), (130, 17)])
@pytest.mark.parametrize("scale", [0.05, 0.6, 3.0])
def test_random_soups_vs_ref_C(kal, O, ref_C, Hh, Ww, scale):
    """ragged image sizes (not multiples of the 16x16 tile), tiny to screen-filling faces (the latter
    exercise the 'large face' lists), valid_faces masks, knum truncation (knum=5)."""
    if Hh * Ww * 3 < 512:
        pytest.skip("reference backward launches 0 blocks when B*H*W < 512 (rasterization_cuda.cu:422-424)")
    d = random_soup(3, 300, scale, seed=Hh * 1000 + Ww)
    d.update(H=Hh, W=Ww)
    full_pipeline_check(kal, O, ref_C, d, knum=5, boxlen=0.05)


def _adversarial_faces(kind, Hh, Ww, seed):
    """Face sets aimed at the span rasterizer's conservative column bounds (dibr_raster_fwd.cuh): pixel centres
    exactly on edges, slivers far thinner than a pixel, zero-area and huge faces."""
    g = torch.Generator().manual_seed(seed)
    if kind == "lattice":
        # vertices ON pixel centres, every 3 pixels: horizontal, vertical and diagonal edges through pixel centres
        xs = (2 * torch.arange(0, Ww, 3, dtype=torch.float32) + 1 - Ww) / Ww
        ys = (Hh - 2 * torch.arange(0, Hh, 3, dtype=torch.float32) - 1) / Hh
        gy, gx = torch.meshgrid(ys, xs, indexing="ij")
        P = torch.stack([gx, gy], -1)
        a, b, c, d = P[:-1, :-1], P[:-1, 1:], P[1:, :-1], P[1:, 1:]
        img = torch.cat([torch.stack([a, b, c], -2).reshape(-1, 3, 2), torch.stack([b, d, c], -2).reshape(-1, 3, 2),
                         torch.stack([a, c, b], -2).reshape(-1, 3, 2)], 0)       # both windings
    elif kind == "slivers":
        n = 600
        ctr = torch.rand(n, 1, 2, generator=g) * 2 - 1
        ang = torch.rand(n, generator=g) * 2 * math.pi
        d = torch.stack([torch.cos(ang), torch.sin(ang)], -1)
        nrm = torch.stack([-d[:, 1], d[:, 0]], -1)
        length = 0.05 + 1.5 * torch.rand(n, 1, generator=g)
        width = 10.0 ** (-6 + 5 * torch.rand(n, 1, generator=g))                # 1e-6 .. 1e-1 NDC
        width[::7] = 0.0                                                        # exactly collinear
        img = ctr + torch.stack([-0.5 * length * d, 0.5 * length * d, (torch.rand(n, 1, generator=g) - 0.5) * length * d
                                 + width * nrm], 1)
    elif kind == "huge":
        n = 40
        img = (torch.rand(n, 3, 2, generator=g) - 0.5) * 10.0 ** (torch.randint(0, 6, (n, 1, 1), generator=g).float())
        img[::5] *= 1e4
    elif kind == "on_vertex":
        # random faces snapped so that one vertex sits on a pixel centre and one edge midpoint on another
        n = 300
        ix = torch.randint(0, Ww, (n,), generator=g).float(); iy = torch.randint(0, Hh, (n,), generator=g).float()
        v0 = torch.stack([(2 * ix + 1 - Ww) / Ww, (Hh - 2 * iy - 1) / Hh], -1)
        jx = torch.randint(0, Ww, (n,), generator=g).float(); jy = torch.randint(0, Hh, (n,), generator=g).float()
        mid = torch.stack([(2 * jx + 1 - Ww) / Ww, (Hh - 2 * jy - 1) / Hh], -1)
        v1 = 2 * mid - v0                                                        # pixel (jx, jy) is the midpoint of v0 v1
        v2 = torch.rand(n, 2, generator=g) * 2 - 1
        img = torch.stack([v0, v1, v2], 1)
    F = img.shape[0]
    z = -1.0 - 2 * torch.rand(1, F, 3, generator=g)
    feat = torch.rand(1, F, 3, 3, generator=g)
    return z.cuda(), img[None].contiguous().cuda(), feat.cuda()


@pytest.mark.parametrize("kind", ["lattice", "slivers", "huge", "on_vertex"])
@pytest.mark.parametrize("Hh,Ww", [(64, 64), (48, 80), (33, 21)])
@pytest.mark.parametrize("multiplier,eps", [(1000., 1e-8), (1., 1e-8), (37.5, 0.), (1000., 1e-3)])
def test_span_rasterizer_adversarial_vs_ref_C(kal, O, ref_C, kind, Hh, Ww, multiplier, eps):
    """face_idx bit-exact and features 1e-5 against the reference's kernel where the conservative spans are tightest."""
    z, img, feat = _adversarial_faces(kind, Hh, Ww, seed=Hh * 100 + Ww)
    out, idx = kal.render.mesh.rasterize(Hh, Ww, z, img, feat, None, multiplier, eps)
    r_feat, r_idx, r_w = O.rasterize_forward(Hh, Ww, z, img, feat, None, multiplier, eps, backend=ref_C)
    assert torch.equal(idx, r_idx), f"face_idx mismatch on {(idx != r_idx).sum().item()} pixels"
    assert (idx >= 0).any() or kind == "huge"
    ok = torch.isfinite(r_feat)
    rel_close(torch.where(ok, out, 0), torch.where(ok, r_feat, 0), 1e-5, "features")


# ----------------------------------------------------------------------------- vs the C++ oracle
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_random_soup_vs_oracle_bit_exact(kal, O, dtype):
    """face_idx AND barycentric weights bit-exact against the CPU restatement (fp32 and fp64)."""
    d = random_soup(2, 500, 0.4, seed=5, dtype=dtype)
    Hh, Ww = 48, 40
    z, img, feat = d["face_vertices_z"], d["face_vertices_image"], d["face_features"][0]
    valid = d["face_normals_z"] >= 0
    o_feat, o_idx, o_w = O.rasterize_forward(Hh, Ww, z.cpu(), img.cpu(), feat.cpu(), valid.cpu())
    img_m = (img * 1000).contiguous()
    # reference-shaped packed entry point (valid faces packed by the caller, as RasterizeCuda.forward does)
    vi = torch.where(valid)
    first = torch.zeros(3, dtype=torch.long, device="cuda")
    first[1:] = torch.cumsum(valid.sum(1), 0)
    v_img = img_m[vi[0], vi[1]].contiguous()
    bb = torch.cat([v_img.min(1)[0], v_img.max(1)[0]], 1).contiguous()
    p_feat, p_sel, p_w = kal._C.render.mesh.packed_rasterize_forward_cuda(
        Hh, Ww, z[vi[0], vi[1]].contiguous(), v_img, bb, feat[vi[0], vi[1]].contiguous(), first, 1000, 1e-8)
    f_feat, f_idx = kal.render.mesh.rasterize(Hh, Ww, z, img, feat, valid)
    assert torch.equal(f_idx.cpu(), o_idx)
    assert torch.equal(p_w.cpu(), o_w), "barycentric weights not bit-exact"
    assert torch.equal(f_feat.cpu(), o_feat) and torch.equal(p_feat.cpu(), o_feat)
    # packed-local index -> original index (rasterization.py:357-363)
    mapped = vi[1][(p_sel + first[:-1].reshape(-1, 1, 1)).reshape(-1).clamp(min=0)].reshape(p_sel.shape)
    mapped[p_sel == -1] = -1
    assert torch.equal(mapped, f_idx)


def test_small_image_backward_works(kal, O):
    """B*H*W < 512: the reference launch fails (floor-division grid); the replacement simply works."""
    d = random_soup(1, 50, 0.8, seed=9)
    z, img, feat = d["face_vertices_z"], d["face_vertices_image"].requires_grad_(True), d["face_features"][0]
    out, idx = kal.render.mesh.rasterize(8, 8, z, img, feat)
    g = torch.rand_like(out)
    out.backward(g)
    o_feat, o_idx, o_w = O.rasterize_forward(8, 8, z.cpu(), img.detach().cpu(), feat.cpu())
    o_gi, o_gf = O.rasterize_backward(g.cpu(), o_idx, o_w, img.detach().cpu(), feat.cpu())
    assert torch.equal(idx.cpu(), o_idx)
    rel_close(img.grad.cpu(), o_gi, 1e-4, "grad img")


def test_edge_cases(kal, O):
    dev = "cuda"
    # no faces at all
    z = torch.zeros(2, 0, 3, device=dev); img = torch.zeros(2, 0, 3, 2, device=dev); feat = torch.zeros(2, 0, 3, 3, device=dev)
    out, idx = kal.render.mesh.rasterize(20, 30, z, img, feat)
    assert out.shape == (2, 20, 30, 3) and torch.all(out == 0) and torch.all(idx == -1)
    soft = kal.render.mesh.dibr_soft_mask(img, idx)
    assert torch.all(soft == 0)
    # all faces invalid
    d = random_soup(2, 40, 0.5, seed=3)
    out, idx = kal.render.mesh.rasterize(33, 33, d["face_vertices_z"], d["face_vertices_image"], d["face_features"][0],
                                         torch.zeros(2, 40, dtype=torch.bool, device=dev))
    assert torch.all(idx == -1) and torch.all(out == 0)
    # degenerate (zero-area) and duplicated faces: lowest index wins exact depth ties
    img = torch.tensor([[[[-.5, -.5], [.5, -.5], [0., .5]]]], device=dev).repeat(1, 4, 1, 1)
    img[0, 3] = 0.25                                          # zero-area face
    z = torch.full((1, 4, 3), -2., device=dev)
    feat = torch.arange(4, device=dev, dtype=torch.float32).view(1, 4, 1, 1).expand(1, 4, 3, 1).contiguous()
    out, idx = kal.render.mesh.rasterize(32, 32, z, img, feat)
    o_feat, o_idx, _ = O.rasterize_forward(32, 32, z.cpu(), img.cpu(), feat.cpu())
    assert torch.equal(idx.cpu(), o_idx) and int(idx.max()) == 0
    # NaN / inf depths and coordinates follow the reference's in-order semantics
    d = random_soup(1, 60, 0.7, seed=11)
    z = d["face_vertices_z"].clone(); img = d["face_vertices_image"].clone()
    z[0, 7, 1] = float("nan"); z[0, 20] = float("inf"); z[0, 31] = -float("inf"); img[0, 44, 2, 0] = float("nan")
    out, idx = kal.render.mesh.rasterize(40, 40, z, img, d["face_features"][0])
    o_feat, o_idx, _ = O.rasterize_forward(40, 40, z.cpu(), img.cpu(), d["face_features"][0].cpu())
    assert torch.equal(idx.cpu(), o_idx)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_nan_depth_in_order_rule(kal, O, ref_C, dtype):
    """rasterization_cuda.cu:162 `if (z0 <= max_z0) continue;` with NaN depths: a NaN-depth face is always accepted,
    the next inside face after it too, then the normal compare resumes.  Stacked identical triangles with NaN
    depths at chosen / random positions, against the oracle's in-order scan and the reference's kernel."""
    tri = torch.tensor([[-.6, -.6], [.6, -.6], [0., .7]])
    # face 0 NaN, face 1 z=-1, face 2 z=-2: the reference keeps face 1 (face 2 is behind it)
    cases = [[float("nan"), -1., -2.], [float("nan"), -2., -1.], [-1., float("nan")], [-1., float("nan"), -5., -3., -4.],
             [float("nan"), float("-inf"), float("-inf")], [-1., float("nan"), float("nan")]]
    g = torch.Generator().manual_seed(3)
    for _ in range(6):
        zz = (-1 - 3 * torch.rand(40, generator=g))
        zz[torch.randint(0, 40, (4,), generator=g)] = float("nan")
        cases.append(zz.tolist())
    for zs in cases:
        F = len(zs)
        img = tri.view(1, 1, 3, 2).repeat(1, F, 1, 1)
        img = (img + 0.02 * torch.rand(1, F, 3, 2, generator=g)).to("cuda", dtype)      # not all pixels see all faces
        z = torch.tensor(zs).view(1, F, 1).repeat(1, 1, 3).to("cuda", dtype)
        feat = torch.rand(1, F, 3, 2, generator=g).to("cuda", dtype)
        out, idx = kal.render.mesh.rasterize(40, 36, z, img, feat)
        o_feat, o_idx, _ = O.rasterize_forward(40, 36, z.cpu(), img.cpu(), feat.cpu())
        r_feat, r_idx, _ = O.rasterize_forward(40, 36, z, img, feat, None, 1000., 1e-8, backend=ref_C)
        assert torch.equal(o_idx, r_idx.cpu()), "oracle vs reference kernel"
        assert torch.equal(idx.cpu(), o_idx), f"depths {zs[:6]}: {(idx.cpu() != o_idx).sum().item()} pixels differ"
    # the first case, spelled out
    img = tri.view(1, 1, 3, 2).repeat(1, 3, 1, 1).to("cuda", dtype)
    z = torch.tensor([float("nan"), -1., -2.]).view(1, 3, 1).repeat(1, 1, 3).to("cuda", dtype)
    _, idx = kal.render.mesh.rasterize(32, 32, z, img, torch.ones(1, 3, 3, 1, device="cuda", dtype=dtype))
    assert int(idx.max()) == 1 and set(idx.unique().tolist()) == {-1, 1}


def test_error_behaviour(kal):
    z = torch.zeros(1, 2, 3); img = torch.zeros(1, 2, 3, 2); feat = torch.zeros(1, 2, 3, 1)
    with pytest.raises(RuntimeError):
        kal.render.mesh.rasterize(8, 8, z, img, feat)                    # CPU tensors -> RuntimeError
    with pytest.raises(ValueError):
        kal.render.mesh.rasterize(8, 8, z.cuda(), img.cuda(), feat.cuda(), backend="opengl")
    with pytest.raises(ValueError):
        kal.render.mesh.rasterize(8, 8, z.cuda(), img.cuda(), feat.cuda(), backend="nvdiffrast")
    with pytest.raises(RuntimeError):
        kal.render.mesh.rasterize(8, 8, z.cuda().half(), img.cuda().half(), feat.cuda().half())
    with pytest.raises(RuntimeError):
        kal._C.render.mesh.dibr_soft_mask_forward_cuda(img.cuda()[:, :, :, :1], torch.zeros(1, 2, 4).cuda(),
                                                        torch.zeros(1, 8, 8, dtype=torch.long).cuda(), 7000, 30, 1000)
    with pytest.warns(UserWarning):
        try:
            kal.render.mesh.rasterize(8, 8, z.cuda(), img.cuda(), feat.cuda(), multiplier=100, backend="nvdiffrast")
        except ValueError:
            pass


def test_knum_truncation_dense(kal, O, ref_C):
    """many faces around uncovered pixels: the first `knum` faces BY INDEX are kept (order semantics)."""
    from kaolin_b200 import synth
    d = to_dev(synth.make_inputs("c2", batch=1))
    d["face_vertices_image"] = (d["face_vertices_image"] * 0.5).contiguous()     # crowd the centre, empty border
    for knum in (1, 7, 30):
        full_pipeline_check(kal, O, ref_C, d, knum=knum, boxlen=0.05)


# ----------------------------------------------------------------------------- binning levels / dense tiles
@pytest.mark.parametrize("Hh,Ww,scale", [(200, 328, 0.25), (320, 200, 0.9), (600, 600, 2.5)])
def test_binning_levels_vs_ref_C(kal, O, ref_C, Hh, Ww, scale):
    """Faces of 40..700 pixels on images of several hundred pixels: fine tile lists, coarse (4x4-tile) cell lists
    and the per-view large list are all populated; enlarged soft-mask boxes push more faces up one level."""
    d = random_soup(2, 160, scale, seed=Hh + Ww)
    d.update(H=Hh, W=Ww)
    full_pipeline_check(kal, O, ref_C, d, knum=12, boxlen=0.04)


def test_dense_tile_multi_pass_vs_ref_C(kal, O, ref_C):
    """More than 2048 candidate faces in one 16x16 tile (config-5 regime): the soft-mask kernel consumes them in
    several face-index ranges; k-lists (knum larger than the hits of some pixels, smaller than others), side
    buffers and gradients must match the reference's single in-order scan."""
    g = torch.Generator().manual_seed(99)
    F = 6000
    ctr = torch.rand(1, F, 1, 2, generator=g) * 0.9 - 0.45
    img = ctr + (torch.rand(1, F, 3, 2, generator=g) - 0.5) * 0.12
    img[:, :, :, 0] += 0.35                                     # leave the left part of the image uncovered
    z = -1.0 - 2 * torch.rand(1, F, 3, generator=g)
    feat = torch.rand(1, F, 3, 2, generator=g)
    d = dict(face_vertices_z=z.cuda(), face_vertices_image=img.cuda(), face_features=[feat.cuda()],
             face_normals_z=None, H=40, W=40)
    for knum in (30, 3000):
        full_pipeline_check(kal, O, ref_C, d, knum=knum, boxlen=0.3)


def test_unused_outputs_cost_no_gradient(kal):
    """Only one of the outputs feeds the loss: the other's gradient is None (no zero-filled tensors), results equal
    the both-outputs run with a zero upstream gradient."""
    d = random_soup(2, 200, 0.5, seed=3)
    z, img0, feat0, nz = d["face_vertices_z"], d["face_vertices_image"], d["face_features"][0], d["face_normals_z"]
    img = img0.clone().requires_grad_(True)
    feat = feat0.clone().requires_grad_(True)
    out, soft, idx = kal.render.mesh.dibr_rasterization(48, 56, z, img, feat, nz)
    soft.sum().backward()
    assert feat.grad is None and torch.isfinite(img.grad).all() and img.grad.abs().sum() > 0
    g_soft_only = img.grad.clone()
    img2 = img0.clone().requires_grad_(True)
    feat2 = feat0.clone().requires_grad_(True)
    out2, soft2, _ = kal.render.mesh.dibr_rasterization(48, 56, z, img2, feat2, nz)
    (soft2.sum() + (out2 * 0).sum()).backward()
    rel_close(img2.grad, g_soft_only, 1e-4, "soft-only gradient")        # atomics: summation order differs run to run
    assert feat2.grad is not None and torch.all(feat2.grad == 0)
    img3 = img0.clone().requires_grad_(True)
    (o_a, o_b), _ = kal.render.mesh.rasterize(48, 56, z, img3, [feat0[..., :1].contiguous().requires_grad_(True),
                                                              feat0[..., 1:].contiguous()], nz >= 0)
    o_a.sum().backward()
    assert torch.isfinite(img3.grad).all()


def test_cuda_graph_capture_matches_eager(kal):
    """The entry points only enqueue work on the caller's stream, so forward + backward can be captured in a CUDA
    graph (kaolin_b200.graphs.GraphedStep) and replayed on new input values in the static buffers."""
    from kaolin_b200.graphs import GraphedStep
    d = random_soup(2, 400, 0.3, seed=11)
    z, nz = d["face_vertices_z"], d["face_normals_z"]
    img_buf = d["face_vertices_image"].clone()
    feat_buf = d["face_features"][0].clone()
    g_out = torch.rand(2, 64, 72, 4, device="cuda")
    g_soft = torch.rand(2, 64, 72, device="cuda")

    def step():
        img = img_buf.detach().requires_grad_(True)
        feat = feat_buf.detach().requires_grad_(True)
        out, soft, idx = kal.render.mesh.dibr_rasterization(64, 72, z, img, feat, nz)
        torch.autograd.backward([out, soft], [g_out, g_soft])
        return out, soft, idx, img.grad, feat.grad

    graphed = GraphedStep(step)
    for shift in (0.0, 0.05):
        img_buf.copy_(d["face_vertices_image"] + shift)
        outs_g = [t.clone() for t in graphed()]
        outs_e = step()
        assert torch.equal(outs_g[2], outs_e[2])
        assert torch.equal(outs_g[0], outs_e[0]) and torch.equal(outs_g[1], outs_e[1])
        for a, b in zip(outs_g[3:], outs_e[3:]):
            scale = max(1.0, float(b.abs().max()))
            assert float((a - b).abs().max()) <= 1e-4 * scale


# ----------------------------------------------------------------------------- full BASELINE sizes, by properties
def _full_size_properties(kal, name, batch, shard):
    """Size-independent properties at a BASELINE.json configuration's full size (the brute-force reference
    cannot run there): run-to-run determinism of the forward, composition identity (dibr_rasterization ==
    rasterize + dibr_soft_mask, test_dibr.py:527-529), view-shard invariance (rendering a contiguous shard of the
    views alone gives exactly the shard of the full-batch result — the multi-GPU partition of SURVEY.md §8e),
    value ranges, and finite gradients that vanish on culled faces."""
    from kaolin_b200 import synth
    d = to_dev(synth.make_inputs(name, batch=batch))
    Hh, Ww = d["H"], d["W"]
    z, img, feat, nz = d["face_vertices_z"], d["face_vertices_image"], d["face_features"][0], d["face_normals_z"]
    img_l = img.clone().requires_grad_(True)
    feat_l = feat.clone().requires_grad_(True)
    out, soft, idx = kal.render.mesh.dibr_rasterization(Hh, Ww, z, img_l, feat_l, nz)
    out2, idx2 = kal.render.mesh.rasterize(Hh, Ww, z, img, feat, nz >= 0)
    soft2 = kal.render.mesh.dibr_soft_mask(img, idx2)
    assert torch.equal(idx, idx2) and torch.equal(out, out2) and torch.equal(soft, soft2)
    covered = idx >= 0
    assert int(idx.max()) < z.shape[1] and int(idx.min()) >= -1
    assert covered.float().mean() > 0.2, "the mesh should cover a good part of the image"
    assert torch.all(soft[covered] == 1) and float(soft.min()) >= 0 and float(soft.max()) <= 1
    assert torch.all(out[~covered] == 0)
    # every rendered face is front-facing (valid_faces = normals_z >= 0, dibr.py:195)
    b_idx = torch.arange(z.shape[0], device="cuda").view(-1, 1, 1).expand_as(idx)
    assert torch.all(nz[b_idx[covered], idx[covered]] >= 0)
    lo, hi = shard
    o_s, s_s, i_s = kal.render.mesh.dibr_rasterization(Hh, Ww, z[lo:hi].contiguous(), img[lo:hi].contiguous(),
                                                       feat[lo:hi].contiguous(), nz[lo:hi].contiguous())
    assert torch.equal(i_s, idx[lo:hi]) and torch.equal(o_s, out[lo:hi]) and torch.equal(s_s, soft[lo:hi])
    g = torch.Generator().manual_seed(5)
    g_out = torch.rand(out.shape[1:], generator=g).cuda().expand_as(out)
    g_soft = torch.rand(soft.shape[1:], generator=g).cuda().expand_as(soft)
    torch.autograd.backward([out, soft], [g_out, g_soft])
    assert torch.isfinite(img_l.grad).all() and torch.isfinite(feat_l.grad).all()
    assert torch.all(feat_l.grad[nz < 0] == 0), "culled faces receive no feature gradient"
    assert feat_l.grad.abs().sum() > 0 and img_l.grad.abs().sum() > 0
    # sum over pixels of the interpolation weights = 1 per covered pixel: grad of a constant-1 upstream w.r.t. the
    # features sums to the number of covered pixels per channel
    feat_c = feat.clone().requires_grad_(True)
    o3, _ = kal.render.mesh.rasterize(Hh, Ww, z, img, feat_c, nz >= 0)
    o3.sum().backward()
    total = feat_c.grad.double().sum().item()
    expect = float(covered.sum().item()) * feat.shape[-1]
    assert abs(total - expect) <= 1e-4 * expect, (total, expect)


def test_c3_full_size_properties(kal):
    """config 3: 100k faces, 64 views, 1024x1024."""
    _full_size_properties(kal, "c3", None, (24, 32))


def test_c5_full_size_properties(kal):
    """config 5 (stress): 1M faces, 8 views, 2048x2048 — coarse-cell lists and multi-pass dense tiles."""
    _full_size_properties(kal, "c5", None, (3, 4))


def test_c4_shape_two_shards(kal):
    """config 4 shape (40k faces, 512x512) on 16 of the 128 cameras."""
    _full_size_properties(kal, "c4", 16, (8, 16))


def test_soft_mask_pair_log_overflow_falls_back_to_recompute(tmp_path):
    """The soft-mask backward normally consumes the forward's pair log; with a log far too small for the scene
    (DIBR_SOFT_LOG_CAP, read once per process -> subprocess) most tiles overflow and are recomputed.  Both must
    give the gradient of the reference's kernels."""
    import os
    import subprocess
    import sys
    script = tmp_path / "run.py"
    script.write_text('''
import sys, torch
sys.path.insert(0, %r)
import kaolin_b200 as kal
from kaolin_b200 import synth
d = synth.make_inputs("c2", batch=2)
img = d["face_vertices_image"].cuda().requires_grad_(True)
z = d["face_vertices_z"].cuda()
feats = d["face_features"][0].cuda()
out, idx = kal.render.mesh.rasterize(d["H"], d["W"], z, img, feats, None, 1000., 1e-8)
soft = kal.render.mesh.dibr_soft_mask(img, idx, d["sigmainv"], d["boxlen"], d["knum"], 1000.)
g = torch.Generator().manual_seed(1)
soft.backward(torch.rand(soft.shape, generator=g).cuda())
torch.save(img.grad.cpu(), sys.argv[1])
''' % os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    grads = []
    for cap in ("0", "20000"):
        env = dict(os.environ, DIBR_B200_TUNING="1", DIBR_SOFT_LOG_CAP=cap)
        out = tmp_path / f"g{cap}.pt"
        subprocess.run([sys.executable, str(script), str(out)], check=True, env=env, timeout=300)
        grads.append(torch.load(out))
    assert float(grads[0].abs().max()) > 0
    rel_close(grads[1], grads[0], 1e-4, "overflowed log vs full log")


@pytest.mark.parametrize("knobs", [
    {"DIBR_RASTER_NT": "128", "DIBR_SOFT_NT": "64"},
    {"DIBR_RASTER_NT": "256", "DIBR_SOFT_NT": "256"},
    {"DIBR_SOFT_SEL": "32"},                                  # many selection passes per soft-mask tile
])
def test_kernel_variants_give_identical_results(tmp_path, knobs):
    """The CTA-size instantiations (64 / 128 / 256 threads per tile) and the candidate-selection target are tuning
    knobs read once per process: every setting must reproduce the default build's face_idx exactly and its soft
    mask / features / gradients within the path's tolerances (the defaults are checked against the reference's
    kernels by every other test of this file)."""
    import os
    import subprocess
    import sys
    script = tmp_path / "run.py"
    script.write_text('''
import sys, torch
sys.path.insert(0, %r)
import kaolin_b200 as kal
from kaolin_b200 import synth
d = synth.make_inputs("c3", batch=1)
img = d["face_vertices_image"].cuda().requires_grad_(True)
z = d["face_vertices_z"].cuda()
feats = d["face_features"][0].cuda().requires_grad_(True)
out, soft, idx = kal.render.mesh.dibr_rasterization(d["H"], d["W"], z, img, feats, d["face_normals_z"].cuda(),
                                                     d["sigmainv"], d["boxlen"], d["knum"])
g = torch.Generator().manual_seed(1)
(out * torch.rand(out.shape, generator=g).cuda()).sum().backward(retain_graph=True)
gi_r = img.grad.clone(); img.grad = None
soft.backward(torch.rand(soft.shape, generator=g).cuda())
torch.save({"idx": idx.cpu(), "out": out.detach().cpu(), "soft": soft.detach().cpu(), "gi_r": gi_r.cpu(),
            "gi_s": img.grad.cpu(), "gf": feats.grad.cpu()}, sys.argv[1])
''' % os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    res = []
    for name, extra in (("default", {}), ("variant", knobs)):
        env = {k: v for k, v in os.environ.items() if not k.startswith("DIBR_")}
        env.update(extra)
        env["DIBR_B200_TUNING"] = "1"            # the one switch that makes the library read its tuning knobs
        out = tmp_path / f"{name}.pt"
        subprocess.run([sys.executable, str(script), str(out)], check=True, env=env, timeout=600)
        res.append(torch.load(out))
    a, b = res
    assert torch.equal(a["idx"], b["idx"])
    assert torch.equal(a["out"], b["out"])
    rel_close(b["soft"], a["soft"], 1e-6, "soft")
    for k in ("gi_r", "gi_s", "gf"):
        rel_close(b[k], a[k], 1e-4, k)
