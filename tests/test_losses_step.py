"""The loss / optimiser side of the multi-view fitting step (BASELINE.json config 4; the reference's training step,
examples/tutorial/ian_dibr.py:214-291): the fused image-L1 loss against its torch composition
(``torch.mean(torch.abs(image - gt_image))``, ian_dibr.py:260), and the CUDA-graph replay of the whole step against
eager launches."""
import math

import pytest
import torch

from helpers import norm_close

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dname", ["f32", "f64"])
@pytest.mark.parametrize("shape", [(3, 17, 23, 3), (1, 1, 1, 1), (2, 128, 128, 3), (70001,)])
def test_image_l1_matches_torch(dname, shape):
    from kaolin_b200.metrics.render import image_l1
    dtype = torch.float32 if dname == "f32" else torch.float64
    g = torch.Generator().manual_seed(3)
    a = torch.rand(shape, generator=g, dtype=torch.float64).to(dtype).cuda()
    b = torch.rand(shape, generator=g, dtype=torch.float64).to(dtype).cuda()
    b.view(-1)[::7] = a.view(-1)[::7]                      # exact zeros: sign(0) = 0 in torch.abs' backward
    res = []
    for fn in (lambda x, y: torch.mean(torch.abs(x - y)), image_l1):
        x, y = a.clone().requires_grad_(True), b.clone().requires_grad_(True)
        loss = fn(x, y)
        (loss * 3.0).backward()
        res.append((loss.detach(), x.grad, y.grad))
    tol = 1e-6 if dname == "f32" else 1e-12
    assert abs(float(res[0][0]) - float(res[1][0])) <= tol * max(1.0, abs(float(res[0][0])))
    assert torch.allclose(res[1][1], res[0][1], rtol=tol, atol=0) and torch.allclose(res[1][2], res[0][2], rtol=tol, atol=0)
    # deterministic: the same bits every time
    assert torch.equal(image_l1(a, b), image_l1(a, b))


def test_image_l1_edge_cases():
    from kaolin_b200.metrics.render import image_l1
    e = torch.zeros(0, 4, 4, 3, device="cuda")
    assert torch.isnan(image_l1(e, e))                     # torch.mean of an empty tensor
    with pytest.raises(RuntimeError):
        image_l1(torch.zeros(2, 2), torch.zeros(2, 2))
    a = torch.rand(4, 8, device="cuda", requires_grad=True)
    image_l1(a, torch.rand(4, 8, device="cuda")).backward()
    assert a.grad.abs().max() <= 1.0 / 32 + 1e-9


def _optimizer(seed=0, cams=4, res=96):
    from kaolin_b200 import multiview as mv
    from kaolin_b200 import synth
    v, f = synth.uv_sphere(48, 24, 0.05, torch.Generator().manual_seed(21))
    v = v / v.norm(dim=1).max()
    rot, eye = synth.orbit_cameras(cams)
    g = torch.Generator().manual_seed(seed)
    vcol = torch.rand(v.shape[0], 3, generator=g)
    tm = (torch.rand(cams, res, res, generator=g) > 0.5).float()
    tf = torch.rand(cams, res, res, 3, generator=g)
    dev = "cuda"
    return mv.MultiViewOptimizer(v.to(dev), f.to(dev), vcol.to(dev), rot.to(dev), eye.to(dev), tm.to(dev), tf.to(dev),
                                 res, res, cams, lr=1e-3)


def test_captured_step_equals_eager_steps():
    eager, graphed = _optimizer(), _optimizer()
    v0 = graphed.vertices.detach().clone()
    assert graphed.capture(warmup=2), graphed.launch_mode
    # capture() ran warm-up steps: restart both from the same state (vertices and Adam moments)
    with torch.no_grad():
        graphed.vertices.copy_(v0)
    for st in graphed.opt.state.values():          # exp_avg, exp_avg_sq, step, arrivals: all start from zero
        for k, t in st.items():
            if torch.is_tensor(t):
                t.zero_()
    losses_e, losses_g = [], []
    for _ in range(3):
        losses_e.append(float(eager.step()))
        losses_g.append(float(graphed.step()))
    for a, b in zip(losses_e, losses_g):
        assert abs(a - b) <= 1e-5 * max(1.0, abs(a)), (losses_e, losses_g)
    norm_close(graphed.vertices - v0, eager.vertices - v0, 1e-3, "vertex update after 3 steps")
    norm_close(graphed.last_grad, eager.last_grad, 1e-3, "vertices.grad of the last step")
    assert "graph" in graphed.launch_mode and eager.launch_mode == "eager launches"


def test_step_loss_matches_unfused_composition():
    """The fitting step's loss through the fused pieces (from_mesh, mask_iou, image_l1) equals the reference-shaped
    composition prepare_vertices -> dibr_rasterization -> torch losses, and so does vertices.grad."""
    from kaolin_b200.render.mesh import prepare_vertices, dibr_rasterization
    from kaolin_b200.metrics.render import mask_iou
    opt = _optimizer(seed=5)
    loss, _, _ = opt.loss()
    loss.backward()
    g_fused = opt.vertices.grad.clone()
    leaf = opt.vertices.detach().clone().requires_grad_(True)
    cam, img, nrm = prepare_vertices(leaf.unsqueeze(0), opt.faces, opt.proj, opt.rot, opt.trans)
    out, soft, _ = dibr_rasterization(opt.height, opt.width, cam[..., 2], img, opt.face_features, nrm[..., 2],
                                      sigmainv=opt.sigmainv, boxlen=opt.boxlen, knum=opt.knum)
    ref = mask_iou(soft, opt.target_mask) + torch.mean(torch.abs(out - opt.target_feat))
    ref.backward()
    assert abs(float(ref) - float(loss)) <= 1e-6 * max(1.0, abs(float(ref)))
    norm_close(g_fused, leaf.grad, 1e-4, "vertices.grad")


@pytest.mark.parametrize("dname", ["f32", "f64"])
def test_per_element_adam_matches_torch(dname):
    from kaolin_b200.multiview import PerElementAdam
    dtype = torch.float32 if dname == "f32" else torch.float64
    g = torch.Generator().manual_seed(9)
    p0 = torch.rand(1, 7001, 3, generator=g, dtype=torch.float64).to(dtype).cuda()
    a, b = p0.clone().requires_grad_(True), p0.clone().requires_grad_(True)
    oa = torch.optim.Adam([a], lr=3e-3, betas=(0.9, 0.99), eps=1e-8)
    ob = PerElementAdam([b], lr=3e-3, betas=(0.9, 0.99), eps=1e-8)
    for it in range(6):
        grad = (torch.rand(p0.shape, generator=g, dtype=torch.float64) - 0.5).to(dtype).cuda() * (10.0 ** (it - 3))
        a.grad, b.grad = grad.clone(), grad.clone()
        oa.step()
        ob.step()
        tol = 2e-6 if dname == "f32" else 1e-12
        assert torch.allclose(b, a, rtol=tol, atol=tol), (it, float((a - b).abs().max()))
    assert float(ob.state[b]["step"]) == 6.0
