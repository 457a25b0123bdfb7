"""CPU suite: pins oracle/ (the C++ restatement) against the reference's golden vectors and against
outputs of the reference's own pure-torch renderer (tests/golden/*.npz, made by make_golden.py).

Mirrors tests/python/kaolin/render/mesh/test_rasterization.py:136-288 and test_dibr.py:109-394,527-529
of the reference with the oracle in the role of the CUDA op."""
import numpy as np
import pytest
import torch

from helpers import golden
from oracle import oracle as O

H, W = 35, 31
DT = {"f32": torch.float32, "f64": torch.float64}


def scene(dname, flip, batch_size):
    g = golden("scenes")
    k = f"{dname}_flip{flip}_"
    return {n: torch.from_numpy(g[k + n])[:batch_size].contiguous()
            for n in ("face_vertices_z", "face_vertices_image", "valid_faces", "face_uvs")}


def mask_iou(lhs, rhs):
    # kaolin/metrics/render.py:18-40
    b = lhs.shape[0]
    mul = lhs * rhs
    add = lhs + rhs
    up = torch.sum(mul.reshape(b, -1), dim=1)
    down = torch.sum((add - mul).reshape(b, -1), dim=1)
    return 1.0 - torch.mean(up / (down + 1e-10))


@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("batch_size", [1, 3])
@pytest.mark.parametrize("flip", [0, 1])
@pytest.mark.parametrize("with_valid", [0, 1])
def test_rasterize_oracle_vs_naive(dname, batch_size, flip, with_valid):
    sc = scene(dname, flip, batch_size)
    n = golden("naive_raster")
    k = f"{dname}_flip{flip}_valid{with_valid}_"
    feats = [sc["face_uvs"], torch.ones_like(sc["face_uvs"][..., :1])]
    valid = sc["valid_faces"] if with_valid else None
    feat, idx, w = O.rasterize_forward(H, W, sc["face_vertices_z"], sc["face_vertices_image"], feats, valid)
    gt_idx = torch.from_numpy(n[k + "face_idx"].astype(np.int64))[:batch_size]
    gt_feat = torch.from_numpy(n[k + "features"])[:batch_size]
    assert torch.equal(idx, gt_idx)                                  # test_rasterization.py:151
    assert torch.allclose(feat, gt_feat, rtol=1e-5, atol=1e-5)      # :152-156
    grad_out = torch.from_numpy(n[f"{dname}_flip{flip}_grad_out"])[:batch_size].contiguous()
    gi, gf = O.rasterize_backward(grad_out, idx, w, sc["face_vertices_image"], feats)
    assert torch.allclose(gi, torch.from_numpy(n[k + "grad_image"])[:batch_size], rtol=1e-3, atol=1e-2)
    assert torch.allclose(gf[..., :2], torch.from_numpy(n[k + "grad_uvs"])[:batch_size], rtol=1e-3, atol=1e-3)
    assert torch.allclose(gf[..., 2:], torch.from_numpy(n[k + "grad_mask"])[:batch_size], rtol=1e-3, atol=1e-3)


# ---------------------------------------------------------------- simple soft-mask goldens
SIMPLE_IMG = [[[[-0.7, 0.], [0., -0.7], [0., 0.7]],
               [[-0.7, 0.], [0., 0.7], [0., -0.7]],
               [[0., -0.7], [0., 0.7], [0.7, 0.]]],
              [[[-0.7, -0.7], [0.7, -0.7], [-0.7, 0.7]],
               [[-0.7, -0.7], [0.7, -0.7], [-0.7, 0.7]],
               [[-0.7, -0.7], [0.7, -0.7], [-0.7, 0.7]]]]
SIMPLE_Z = [[[-2., -1., -1.], [-2.5, -3., -3.], [-2., -2., -2.]],
            [[-2., -1., -3.], [-2., -2., -2.], [-2., -3., -1.]]]


def simple_inputs(dtype):
    img = torch.tensor(SIMPLE_IMG, dtype=dtype)
    z = torch.tensor(SIMPLE_Z, dtype=dtype)
    feat = torch.zeros(z.shape + (1,), dtype=dtype)
    _, idx, _ = O.rasterize_forward(H, W, z, img, feat)
    return img, idx


@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("sigmainv", [7000, 70])
@pytest.mark.parametrize("boxlen", [0.02, 0.2])
@pytest.mark.parametrize("multiplier", [1000, 100, 1])
@pytest.mark.parametrize("knum", [30, 20])
def test_simple_soft_mask(dname, sigmainv, boxlen, multiplier, knum):
    dtype = DT[dname]
    g = golden("dibr_simple")
    tag = f"{H}_{W}_{int(sigmainv)}_{boxlen}"
    img, idx = simple_inputs(dtype)
    soft, img_m, prob, cidx, ctype = O.dibr_soft_mask_forward(img, idx, sigmainv, boxlen, knum, multiplier)
    gt_soft = torch.from_numpy(g["soft_mask_" + tag]).to(dtype)
    gt_idx = torch.from_numpy(g["close_face_idx_" + tag].astype(np.int64)) - 1
    gt_prob = torch.from_numpy(g["close_face_dist_" + tag]).to(dtype)
    gt_type = torch.from_numpy(g["close_face_dist_type_" + tag])
    assert torch.allclose(soft, gt_soft, atol=1e-5, rtol=1e-5)      # test_dibr.py:133-140
    assert torch.equal(cidx, gt_idx[..., :knum])
    assert torch.allclose(prob, gt_prob[..., :knum], atol=1e-5, rtol=1e-5)
    assert torch.equal(ctype, gt_type[..., :knum])
    # backward through mask_iou — test_dibr.py:163-191
    soft_leaf = soft.clone().requires_grad_(True)
    mask = idx != -1
    shifted = torch.nn.functional.pad(mask, (0, 5))[..., 5:]
    mask_iou(soft_leaf, shifted).backward()
    gi = O.dibr_soft_mask_backward(soft_leaf.grad, soft, img_m, idx, prob, cidx, ctype, sigmainv, multiplier)
    gt_grad = torch.from_numpy(g["grad_face_vertices_image_" + tag]).to(dtype)
    assert torch.allclose(gi, gt_grad, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("batch_size", [1, 3])
@pytest.mark.parametrize("flip", [0, 1])
@pytest.mark.parametrize("sigmainv", [7000, 70])
@pytest.mark.parametrize("boxlen", [0.02, 0.01])
@pytest.mark.parametrize("multiplier,knum", [(1000, 30), (100, 40), (1, 40)])
def test_sphere_soft_mask(dname, batch_size, flip, sigmainv, boxlen, multiplier, knum):
    dtype = DT[dname]
    g = golden("dibr_sphere")
    tag = f"{H}_{W}_{int(sigmainv)}_{boxlen}"
    sc = scene(dname, flip, batch_size)
    img = sc["face_vertices_image"]
    feat = torch.zeros(sc["face_vertices_z"].shape + (1,), dtype=dtype)
    _, idx, _ = O.rasterize_forward(H, W, sc["face_vertices_z"], img, feat)
    soft, img_m, prob, cidx, ctype = O.dibr_soft_mask_forward(img, idx, sigmainv, boxlen, knum, multiplier)
    gt_soft = torch.from_numpy(g["soft_mask_" + tag]).to(dtype)[:batch_size]
    gt_idx = torch.from_numpy(g["close_face_idx_" + tag].astype(np.int64))[:batch_size] - 1
    gt_prob = torch.from_numpy(g["close_face_dist_" + tag]).to(dtype)[:batch_size]
    gt_type = torch.from_numpy(g["close_face_dist_type_" + tag])[:batch_size]
    if multiplier != 1:                     # the reference checks forward values at 1000 and 100 only (:309,344)
        assert torch.allclose(soft, gt_soft, atol=1e-5, rtol=1e-5)  # test_dibr.py:333-341
        assert torch.equal(cidx, gt_idx[..., :knum])
        assert torch.allclose(prob, gt_prob[..., :knum], atol=1e-5, rtol=1e-5)
        diff = ctype != gt_type[..., :knum]
        assert diff.sum().item() / diff.numel() <= 0.01
    soft_leaf = soft.clone().requires_grad_(True)
    shifted = torch.nn.functional.pad(idx != -1, (0, 5))[..., 5:]
    mask_iou(soft_leaf, shifted).backward()
    gi = O.dibr_soft_mask_backward(soft_leaf.grad, soft, img_m, idx, prob, cidx, ctype, sigmainv, multiplier)
    gt_grad = torch.from_numpy(g["grad_face_vertices_image_" + tag]).to(dtype)[:batch_size]
    assert torch.allclose(gi, gt_grad, rtol=1e-1, atol=1e-1)        # test_dibr.py:390-394


def test_eps_add_equivalence():
    """fp32: float(double(n) + copysign(eps, n)) == n + copysignf(eps, n) bit for bit."""
    import ctypes
    rng = np.random.default_rng(7)
    bits = rng.integers(0, 2**32, size=4_000_000, dtype=np.uint64).astype(np.uint32)
    vals = bits.view(np.float32).copy()
    special = np.array([0.0, -0.0, 1e-8, -1e-8, 2e-8, 1e-38, -1e-38, 1e-45, np.inf, -np.inf, 1.0, -1.0,
                        5e-9, 1.5e-8, 2.0**-27, 2.0**-26, 3e-1, 1e6], dtype=np.float32)
    near = (np.float32(1e-8) * rng.uniform(0.25, 4e9, size=1_000_000)).astype(np.float32)
    vals = np.concatenate([vals, special, near, -near])
    for eps in (1e-8, 1e-6, 1e-10, 0.0, 1.0):
        bad = O.lib().oracle_check_eps_add_equivalence(
            ctypes.c_void_p(vals.ctypes.data), ctypes.c_int64(vals.size), ctypes.c_float(eps))
        assert bad == 0
