"""Shared test helpers (no dependency on /root/reference at run time)."""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_obj(path):
    """Minimal Wavefront OBJ reader: triangles with ``v/vt[/vn]`` corners (enough for the reference's
    tests/samples/model.obj).  Returns 0-based numpy arrays."""
    vs, vts, fv, fvt = [], [], [], []
    with open(path) as fh:
        for line in fh:
            t = line.split()
            if not t:
                continue
            if t[0] == "v":
                vs.append([float(x) for x in t[1:4]])
            elif t[0] == "vt":
                vts.append([float(x) for x in t[1:3]])
            elif t[0] == "f":
                corners = [c.split("/") for c in t[1:]]
                assert len(corners) == 3, "triangles only"
                fv.append([int(c[0]) - 1 for c in corners])
                fvt.append([int(c[1]) - 1 if len(c) > 1 and c[1] else 0 for c in corners])
    return dict(vertices=np.asarray(vs, dtype=np.float32), uvs=np.asarray(vts, dtype=np.float32),
                faces=np.asarray(fv, dtype=np.int64), face_uvs_idx=np.asarray(fvt, dtype=np.int64))


_CACHE = {}


def golden(name):
    """Load tests/golden/<name>.npz once."""
    if name not in _CACHE:
        _CACHE[name] = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    return _CACHE[name]


def elem_close(a, b, rel, name, atol=1e-6):
    """|a - b| <= rel * |b| + atol for EVERY element (north_star: "within 1e-5 / 1e-4 relative"); reports the
    worst element.  `atol` covers values that cancel to ~0 (soft = 1 - prod(1 - prob) near 0)."""
    import torch
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (name, tuple(a.shape), tuple(b.shape))
    if b.numel() == 0:
        return
    excess = ((a - b).abs() - (rel * b.abs() + atol)).reshape(-1)
    excess = torch.where(torch.isnan(excess), torch.full_like(excess, float("inf")), excess)
    both_nan = (torch.isnan(a) & torch.isnan(b)).reshape(-1)
    excess = torch.where(both_nan, torch.full_like(excess, -1.0), excess)
    worst = int(excess.argmax())
    idx = tuple(int(i) for i in torch.unravel_index(torch.tensor(worst), b.shape))
    assert float(excess[worst]) <= 0, (f"{name}: worst element {idx}: got {float(a.reshape(-1)[worst]):.9g}, "
                                       f"reference {float(b.reshape(-1)[worst]):.9g} (rel {rel}, atol {atol})")


def norm_close(a, b, rel, name):
    """max|a - b| <= rel * max|b|: for gradients, which the reference accumulates with unordered atomics (its own
    tests use rtol 1e-3 / atol 1e-2, test_dibr.py:351-394)."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    scale = max(1e-30, float(b.abs().max()))
    err = float((a - b).abs().max())
    assert err <= rel * scale, f"{name}: max abs err {err:.3e} > {rel} * {scale:.3e}"
