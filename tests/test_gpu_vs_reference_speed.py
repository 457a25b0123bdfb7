"""The reference's own CUDA kernels (oracle/_ref/ref_C.so: kaolin/csrc/render/mesh/*.cu compiled unmodified for
sm_100a) behind the reference's torch glue, timed on the same B200 next to this library, on BASELINE.json
configs 1 and 2 (rasterize + dibr_soft_mask, forward + backward).  Device-timed with CUDA events.  The table is
written to gpurun_out/vs_reference_cuda.json when that directory exists (DESIGN.md §7 quotes it)."""
import json
import os

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RESULTS = {}


def _time(fn, iters, warm):
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


@pytest.mark.gpu
@pytest.mark.parametrize("name,iters", [("c1", 10), ("c2", 2)])
def test_step_time_vs_reference_cuda_kernels(name, iters):
    import kaolin_b200 as kal
    from kaolin_b200 import synth
    from oracle import oracle as O
    ref_C = O.load_ref_C()
    if ref_C is None:
        pytest.skip("oracle/_ref/ref_C.so not built")
    d = synth.make_inputs(name)
    Hh, Ww = d["H"], d["W"]
    z = d["face_vertices_z"].cuda()
    img = d["face_vertices_image"].cuda()
    feats = torch.cat([f.cuda() for f in d["face_features"]], -1).contiguous()
    valid = None if d["face_normals_z"] is None else (d["face_normals_z"].cuda() >= 0.)
    g = torch.Generator().manual_seed(1)
    B, C = z.shape[0], feats.shape[-1]
    g_feat = torch.rand(B, Hh, Ww, C, generator=g).cuda()
    g_soft = torch.rand(B, Hh, Ww, generator=g).cuda()
    sig, box, knum, mult, eps = d["sigmainv"], d["boxlen"], d["knum"], 1000., 1e-8

    def ours():
        il, fl = img.clone().requires_grad_(True), feats.clone().requires_grad_(True)
        out, idx = kal.render.mesh.rasterize(Hh, Ww, z, il, fl, valid, mult, eps)
        soft = kal.render.mesh.dibr_soft_mask(il, idx, sig, box, knum, mult)
        torch.autograd.backward([out, soft], [g_feat, g_soft])
        return il.grad

    def reference():            # RasterizeCuda / DibrSoftMaskCuda forward+backward of the reference, on ref_C
        r_feat, r_idx, r_w = O.rasterize_forward(Hh, Ww, z, img, feats, valid, mult, eps, backend=ref_C)
        r_soft, r_img_m, r_prob, r_cidx, r_ctype = O.dibr_soft_mask_forward(img, r_idx, sig, box, knum, mult,
                                                                             backend=ref_C)
        gi, gf = O.rasterize_backward(g_feat, r_idx, r_w, img, feats, eps, backend=ref_C,
                                      interpolated_features=r_feat)
        gs = O.dibr_soft_mask_backward(g_soft, r_soft, r_img_m, r_idx, r_prob, r_cidx, r_ctype, sig, mult,
                                       backend=ref_C)
        return gi + gs

    t_ours = _time(ours, iters * 5, 3)
    t_ref = _time(reference, iters, 1)
    RESULTS[name] = {"ours_ms": round(t_ours, 4), "reference_cuda_kernels_ms": round(t_ref, 3),
                     "speedup": round(t_ref / t_ours, 1), "pixels": B * Hh * Ww, "faces": int(z.shape[1])}
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "vs_reference_cuda.json"), "w") as fh:
            json.dump(RESULTS, fh, indent=1)
    assert t_ours < t_ref


@pytest.mark.gpu
def test_deftet_time_vs_reference_cuda_kernels():
    """deftet_sparse_render forward + backward: 100k faces, 4 x 16384 query points, knum 300 (the reference's
    default), against DeftetSparseRenderer's torch glue on the reference's own kernels."""
    import kaolin_b200 as kal
    from oracle import oracle as O
    ref_C = O.load_ref_C()
    if ref_C is None or not hasattr(ref_C.render.mesh, "deftet_sparse_render_forward_cuda"):
        pytest.skip("oracle/_ref/ref_C.so (with deftet) not built")
    g = torch.Generator().manual_seed(5)
    B, F, P, knum = 4, 100000, 16384, 300
    centre = torch.rand(B, F, 1, 2, generator=g) * 2 - 1
    img = (centre + (torch.rand(B, F, 3, 2, generator=g) - 0.5) * 0.05).cuda()
    z = (-(torch.rand(B, F, 3, generator=g) * 4 + 0.5)).cuda()
    pix = (torch.rand(B, P, 2, generator=g) * 2 - 1).cuda()
    rr = torch.tensor([[[-10., 0.]]]).repeat(B, P, 1).cuda()
    feats = torch.rand(B, F, 3, 4, generator=g).cuda()
    gout = torch.rand(B, P, knum, 4, generator=g).cuda()

    def ours():
        il, fl = img.clone().requires_grad_(True), feats.clone().requires_grad_(True)
        out, idx = kal.render.mesh.deftet_sparse_render(pix, rr, z, il, fl, knum)
        out.backward(gout)
        return idx

    def reference():
        out, idx, w = O.deftet_sparse_render_forward(pix, rr, z, img, feats, knum, 1e-8, backend=ref_C)
        O.deftet_sparse_render_backward(gout, idx, w, img, feats, 1e-8, backend=ref_C)
        return idx

    assert torch.equal(ours(), reference())
    t_ours, t_ref = _time(ours, 10, 2), _time(reference, 2, 1)
    RESULTS["deftet"] = {"ours_ms": round(t_ours, 4), "reference_cuda_kernels_ms": round(t_ref, 3),
                         "speedup": round(t_ref / t_ours, 1), "points": B * P, "faces": F, "knum": knum}
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "vs_reference_cuda.json"), "w") as fh:
            json.dump(RESULTS, fh, indent=1)
    assert t_ours < t_ref
