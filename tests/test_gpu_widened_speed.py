"""Device-timed comparison of the widened rows with the reference's own torch compositions
(the oracle restatements are the reference's torch code, run here on the same GPU): the fused kernels must not be
slower.  The measured table is written to gpurun_out/widened_ops.json when that directory exists (DESIGN.md §9
quotes it).  Sizes follow BASELINE.json config 4 (40k faces, 16 cameras per GPU, 512x512)."""
import json
import math
import os

import pytest
import torch

from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RESULTS = {}


def _time(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def _record(name, ours_ms, ref_ms):
    RESULTS[name] = {"ours_ms": round(ours_ms, 4), "torch_reference_ms": round(ref_ms, 4),
                     "speedup": round(ref_ms / ours_ms, 2)}
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "widened_ops.json"), "w") as fh:
            json.dump(RESULTS, fh, indent=1)


@pytest.mark.gpu
def test_prepare_vertices_speed():
    from kaolin_b200 import synth
    from kaolin_b200.render.mesh import prepare_vertices
    d = synth.make_inputs("c4", views=list(range(16)))
    dev = "cuda"
    v = d["vertices"].to(dev).unsqueeze(0)
    f, rot, eye = d["faces"].to(dev), d["rot"].to(dev), d["eye"].to(dev)
    proj = torch.tensor([1 / math.tan(math.pi / 8), 1 / math.tan(math.pi / 8), -1.0], device=dev)
    B = rot.shape[0]
    ups = None

    def run(fn, expand):
        nonlocal ups
        leaf = v.clone().requires_grad_(True)
        out = fn(leaf.expand(B, -1, -1) if expand else leaf, f, proj, rot, eye)
        if ups is None:
            ups = [torch.randn_like(o) for o in out]
        torch.autograd.backward(list(out), ups)
        return leaf.grad

    ours = _time(lambda: run(prepare_vertices, False))
    ref = _time(lambda: run(O.prepare_vertices, True))
    _record("prepare_vertices fwd+bwd (40k faces x 16 cameras)", ours, ref)
    assert ours < ref


@pytest.mark.gpu
def test_texture_mapping_speed():
    from kaolin_b200.render.mesh import texture_mapping
    g = torch.Generator().manual_seed(3)
    B, H, W = 16, 512, 512
    uv = torch.rand(B, H, W, 2, generator=g).cuda()
    mask = (torch.rand(B, H, W, 1, generator=g) > 0.3).float().cuda()
    tex = torch.rand(1, 3, 512, 512, generator=g).cuda()
    gout = torch.randn(B, H, W, 3, generator=g).cuda()

    def ours_step():
        t = tex.clone().requires_grad_(True)
        texture_mapping(uv, t, mode="bilinear", mask=mask).backward(gout)
        return t.grad

    def ref_step():                                   # examples/tutorial/ian_dibr.py:254-258
        t = tex.clone().requires_grad_(True)
        img = O.texture_mapping(uv, t.repeat(B, 1, 1, 1), mode="bilinear")
        torch.clamp(img * mask, 0., 1.).backward(gout)
        return t.grad

    a, b = ours_step(), ref_step()
    assert float((a - b).abs().max()) <= 1e-4 * float(b.abs().max())
    ours, ref = _time(ours_step), _time(ref_step)
    _record("texture_mapping + clamp(image*mask) fwd+bwd (16 x 512x512, 512x512 texture)", ours, ref)
    assert ours < ref


@pytest.mark.gpu
def test_mask_iou_speed():
    from kaolin_b200.metrics.render import mask_iou
    g = torch.Generator().manual_seed(4)
    lhs = torch.rand(16, 512, 512, generator=g).cuda()
    rhs = (torch.rand(16, 512, 512, generator=g) > 0.5).float().cuda()

    def run(fn):
        leaf = lhs.clone().requires_grad_(True)
        fn(leaf, rhs).backward()
        return leaf.grad

    ours, ref = _time(lambda: run(mask_iou)), _time(lambda: run(O.mask_iou))
    _record("mask_iou fwd+bwd (16 x 512x512)", ours, ref)
    assert ours < ref
