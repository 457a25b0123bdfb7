"""Multi-GPU plumbing of the DIB-R path: one process per GPU, the batch of views is sharded
contiguously across ranks, nothing on the data path is exchanged.  The only collective is ONE sum all-reduce
per optimisation step of the LEAF gradients (vertices, optionally a texture) when the views share
parameters — the multi-view shape optimisation of BASELINE.json config 4 (the reference's training step,
examples/tutorial/ian_dibr.py:214-291, run on one GPU there).

torch.distributed is plumbing here (NCCL over NVLink on the GPU box, gloo in the CPU tests); the rendering
kernels are the C-ABI library.
"""
import math

import torch
import torch.distributed as dist


def shard_range(n_items, rank, world_size):
    """Contiguous shard [lo, hi) of ``n_items`` views for ``rank``: sizes differ by at most one, earlier ranks
    take the remainder (B=64 -> 8 per GPU at 8 ranks; B=128 cameras -> 16 per GPU)."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError(f"invalid rank {rank} for world size {world_size}")
    base, rem = divmod(int(n_items), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def all_reduce_leaf_grads(params, group=None, average=False):
    """Sum (or average) the ``.grad`` of the leaf tensors over all ranks with a SINGLE collective: the grads are
    packed into one flat fp32/fp64 bucket (≈0.24 MB for 20k vertices, +3 MB for a 512² texture — latency
    bound, so one launch), all-reduced on the current stream, and copied back.  Leaves without a grad
    contribute zeros so that every rank issues the same collective.  No-op without an initialised group."""
    params = [p for p in params if p is not None]
    if not params:
        return None
    if len(params) == 1 and params[0].grad is not None and params[0].grad.is_contiguous():
        # a single leaf (the usual case: the vertices): reduce its gradient in place, no packing kernels
        g = params[0].grad
        if dist.is_available() and dist.is_initialized():
            dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)
            if average:
                g /= dist.get_world_size(group)
        return g.view(-1)
    dtype = params[0].dtype
    device = params[0].device
    sizes = [p.numel() for p in params]
    bucket = torch.zeros(sum(sizes), dtype=dtype, device=device)
    off = 0
    for p, n in zip(params, sizes):
        if p.grad is not None:
            bucket[off:off + n].copy_(p.grad.reshape(-1))
        off += n
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(bucket, op=dist.ReduceOp.SUM, group=group)
        if average:
            bucket /= dist.get_world_size(group)
    off = 0
    for p, n in zip(params, sizes):
        g = bucket[off:off + n].view_as(p)
        if p.grad is None:
            p.grad = g.clone()
        else:
            p.grad.copy_(g)
        off += n
    return bucket


def look_at(eye, at, up):
    """Camera rotation / translation of kaolin/render/camera/legacy.py:39-83 (generate_rotate_translate_matrices)."""
    cz = at - eye
    cz = cz / (cz.norm(dim=1, keepdim=True) + 1e-10)
    cx = torch.cross(cz, up.expand_as(cz), dim=1)
    cx = cx / (cx.norm(dim=1, keepdim=True) + 1e-10)
    cy = torch.cross(cx, cz, dim=1)
    cy = cy / (cy.norm(dim=1, keepdim=True) + 1e-10)
    return torch.stack([cx, cy, -cz], dim=1), eye


class PerElementAdam(torch.optim.Optimizer):
    """``torch.optim.Adam(params, lr, betas, eps)`` (no weight decay, no amsgrad — the fitting loop's optimiser,
    examples/tutorial/ian_dibr.py:116-123) with one CUDA thread per element (``dibr_adam_step_*``).  torch's fused /
    foreach Adam gives a tensor to ``multi_tensor_apply`` in chunks of 65 536 elements per CTA, so the 60 k coordinates of
    a 20 k-vertex mesh run on ONE CTA (49 us of a 600 us step).  The step counter lives on the device: capturable."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, closure=None):
        from . import _lib
        lib = _lib.load()
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or not p.is_contiguous() or not p.grad.is_contiguous():
                    raise RuntimeError("PerElementAdam: parameters and gradients must be contiguous CUDA tensors")
                st = self.state[p]
                if not st:
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                    st["step"] = torch.zeros(1, dtype=torch.float32, device=p.device)
                    st["arrivals"] = torch.zeros(1, dtype=torch.int32, device=p.device)
                suf = _lib.suffix(p.dtype, "adam_step")
                with torch.cuda.device(p.device):
                    code = getattr(lib, "dibr_adam_step_" + suf)(
                        _lib.ptr(p), _lib.ptr(p.grad), _lib.ptr(st["exp_avg"]), _lib.ptr(st["exp_avg_sq"]),
                        _lib.ptr(st["step"]), _lib.ptr(st["arrivals"]), p.numel(), float(group["lr"]), float(b1),
                        float(b2), float(group["eps"]), _lib.current_stream(p.device))
                _lib.check(code, "adam_step")
        return None


class MultiViewOptimizer:
    """Multi-view DIB-R shape optimisation (BASELINE.json config 4).  Each rank owns a contiguous shard of the
    cameras; the mesh (vertices leaf, faces, per-vertex features) is replicated.  One step =
    dibr_rasterization_from_mesh (prepare_vertices folded into the setup kernel) -> mask IoU + image L1 loss (fused
    kernels) -> backward -> ONE all-reduce of ``vertices.grad`` -> Adam (one thread per coordinate), replayed from one CUDA graph after
    ``capture()``.  Losses are sums over the local views divided by the GLOBAL view count, so the all-reduced gradient
    equals the single-process gradient over all cameras."""

    def __init__(self, vertices, faces, vertex_features, rot, trans, target_mask, target_feat, height, width,
                 n_views_total, lr=1e-3, sigmainv=7000., boxlen=0.02, knum=30, group=None, fovy=math.pi / 4):
        self.vertices = vertices.detach().clone().requires_grad_(True)
        self.faces = faces
        self.vertex_features = vertex_features
        self.rot, self.trans = rot, trans
        self.target_mask, self.target_feat = target_mask, target_feat
        self.height, self.width = height, width
        self.n_views_total = n_views_total
        self.sigmainv, self.boxlen, self.knum = sigmainv, boxlen, knum
        self.group = group
        # kaolin/render/camera/legacy.py:141-159 (generate_perspective_projection, ratio 1)
        self.proj = torch.tensor([1.0 / math.tan(fovy / 2.0), 1.0 / math.tan(fovy / 2.0), -1.0],
                                 dtype=vertices.dtype, device=vertices.device)
        # the per-face features do not depend on the vertices being optimised: gathered once
        self.face_features = vertex_features[faces].unsqueeze(0).expand(rot.shape[0], -1, -1, -1).contiguous()
        # Adam with one thread per coordinate on the GPU (torch's foreach path: 4-5 single-CTA launches per step; its
        # fused path: one single-CTA launch of 49 us)
        self.opt = PerElementAdam([self.vertices], lr=lr) if self.vertices.is_cuda else \
            torch.optim.Adam([self.vertices], lr=lr)
        self.last_grad = None               # the all-reduced vertices.grad of the last step
        self.launch_mode = "eager launches"
        self.time_all_reduce = False        # measurement: CUDA events around the collective on the compute stream
        self._ar_events = []
        self._graph = None                  # CUDA graph of one whole step (capture())
        self._graph_loss = None

    def all_reduce_ms(self):
        """Durations of the all-reduce calls made while ``time_all_reduce`` was set (synchronises)."""
        torch.cuda.synchronize()
        ms = [a.elapsed_time(b) for a, b in self._ar_events]
        self._ar_events = []
        return ms or [0.0]

    def loss(self):
        from .render.mesh import dibr_rasterization_from_mesh
        from .metrics.render import mask_iou as fused_mask_iou
        verts = self.vertices if self.vertices.dim() == 3 else self.vertices.unsqueeze(0)
        b = self.rot.shape[0]
        # prepare_vertices + dibr_rasterization (ian_dibr.py:246-252) in one native call each way
        img, soft, face_idx = dibr_rasterization_from_mesh(self.height, self.width, verts, self.faces,
                                                           self.face_features, self.proj, camera_rot=self.rot,
                                                           camera_trans=self.trans, sigmainv=self.sigmainv,
                                                           boxlen=self.boxlen, knum=self.knum)
        # per-view IoU and L1 terms, normalised by the global number of views:
        # sum_local(1 - iou_b) / n_total = mask_iou(local) * b / n_total   (fused reduction + analytic backward)
        # (every view has the same number of pixels: the sum of the per-view L1 means is b times the shard's mean)
        from .metrics.render import image_l1
        l_mask = fused_mask_iou(soft, self.target_mask) * (b / self.n_views_total)
        l_feat = image_l1(img, self.target_feat) * (b / self.n_views_total)
        return l_mask + l_feat, soft, face_idx

    def capture(self, warmup=3):
        """Capture one whole step — prepare + render + losses + backward + the NCCL all-reduce + Adam — into a CUDA
        graph; ``step()`` then replays it (one launch per step: the step is ~35 launches of 3-150 us each, so eager
        dispatch costs as much as the kernels).  The library only enqueues work on the current stream, NCCL
        collectives and the fused capturable Adam are capturable.  Camera tensors, targets and ``vertices`` are the
        static inputs (update them in place).  Falls back to eager launches, with the reason in ``launch_mode``."""
        if not self.vertices.is_cuda:
            return False
        dev = self.vertices.device
        try:
            side = torch.cuda.Stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(warmup):
                    self._eager_step()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            self.opt.zero_grad(set_to_none=True)       # grads are allocated from the graph's pool during capture
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self._graph_loss = self._eager_step(zero=False)
            self._graph = graph
            self.launch_mode = "CUDA graph replay of the whole step (render, losses, backward, all-reduce, Adam)"
            return True
        except Exception as e:                          # noqa: BLE001 - any capture failure means: stay eager
            self._graph = None
            torch.cuda.synchronize(dev)
            self.launch_mode = f"eager launches (graph capture failed: {type(e).__name__}: {str(e)[:120]})"
            self.opt.zero_grad(set_to_none=True)
            return False

    def step(self):
        if self._graph is not None and not self.time_all_reduce:
            self._graph.replay()
            self.last_grad = self.vertices.grad
            return self._graph_loss
        return self._eager_step()

    def _eager_step(self, zero=True):
        if zero:
            # after capture() the graph owns vertices.grad: keep that tensor (zero it in place) so that replays and
            # eager measurement steps can be mixed
            self.opt.zero_grad(set_to_none=self._graph is None)
        loss, _, _ = self.loss()
        loss.backward()
        if self.time_all_reduce and self.vertices.is_cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            all_reduce_leaf_grads([self.vertices], group=self.group)
            b.record()
            self._ar_events.append((a, b))
        else:
            all_reduce_leaf_grads([self.vertices], group=self.group)
        self.last_grad = self.vertices.grad
        self.opt.step()
        return loss.detach()
