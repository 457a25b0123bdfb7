"""CPU tests of the multi-GPU host logic with the gloo backend, world_size 2: view sharding and the
single-collective all-reduce of leaf gradients.  The rendering kernels need a GPU, so the per-view loss here
is a differentiable surrogate on top of prepare_vertices; what is checked is the plumbing: the all-reduced
gradient equals the single-process gradient over all views."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from kaolin_b200 import multiview as mv
from kaolin_b200 import synth


def test_shard_range_partitions():
    for n in (0, 1, 7, 64, 128, 129):
        for world in (1, 2, 3, 4, 8):
            spans = [mv.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and a <= b
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    assert mv.shard_range(64, 3, 8) == (24, 32)
    assert mv.shard_range(128, 7, 8) == (112, 128)
    with pytest.raises(ValueError):
        mv.shard_range(4, 2, 2)


def _surrogate_loss(vertices, faces, rot, trans, n_total):
    fcam, fimg, nrm = synth.prepare_vertices(vertices, faces, rot, trans)
    return (fimg.square().sum() + (nrm[..., 2] * fcam[..., 2].mean(-1)).sum()) / n_total


def _scene(n_views=6):
    v, f = synth.icosphere(1)
    v = v.double()
    rot, eye = synth.orbit_cameras(n_views, dtype=torch.float64)
    return v, f, rot, eye


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        v, f, rot, eye = _scene()
        n = rot.shape[0]
        lo, hi = mv.shard_range(n, rank, world)
        vert = v.clone().requires_grad_(True)
        extra = torch.zeros(3, dtype=torch.float64, requires_grad=True)     # a leaf that gets no gradient on rank 1
        loss = _surrogate_loss(vert, f, rot[lo:hi], eye[lo:hi], n)
        if rank == 0:
            loss = loss + extra.sum()
        loss.backward()
        single = vert.detach().clone().requires_grad_(True)          # the one-leaf path reduces in place
        single.grad = vert.grad.clone()
        bucket = mv.all_reduce_leaf_grads([vert, extra])
        assert bucket.numel() == vert.numel() + 3
        flat = mv.all_reduce_leaf_grads([single])
        assert flat.data_ptr() == single.grad.data_ptr() and torch.equal(single.grad, vert.grad)
        torch.save({"vert": vert.grad.clone(), "extra": extra.grad.clone()}, os.path.join(out_dir, f"g{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_all_reduce_leaf_grads_gloo_world2(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    v, f, rot, eye = _scene()
    vert = v.clone().requires_grad_(True)
    _surrogate_loss(vert, f, rot, eye, rot.shape[0]).backward()
    for r in range(2):
        g = torch.load(os.path.join(str(tmp_path), f"g{r}.pt"))
        assert torch.allclose(g["vert"], vert.grad, rtol=1e-12, atol=1e-14)
        assert torch.equal(g["extra"], torch.ones(3, dtype=torch.float64))


def test_all_reduce_without_group_is_identity():
    p = torch.ones(4, requires_grad=True)
    (p * 3).sum().backward()
    mv.all_reduce_leaf_grads([p, None])
    assert torch.equal(p.grad, torch.full((4,), 3.0))
