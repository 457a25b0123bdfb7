"""Golden vectors of the reference's ``prepare_vertices`` (kaolin/render/mesh/utils.py:128-175) — run ONCE in the
build container, where /root/reference exists:

    python tests/golden/make_golden_prepare.py        ->  tests/golden/prepare_vertices.npz

The reference modules are pure torch here (camera/legacy.py, ops/mesh/mesh.py, ops/mesh/trianglemesh.py,
render/mesh/utils.py); they are loaded BY PATH under stub packages (no kaolin._C needed) and run on the CPU:
outputs and autograd gradients for fp32 / fp64, with (camera_rot, camera_trans) and with camera_transform,
vertices batched (B,V,3) or shared (1,V,3 expanded).  Nothing reads /root/reference at test time.
"""
import importlib.util
import math
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("DIBR_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, os.pardir))
from helpers import load_obj  # noqa: E402


def _pkg(name):
    m = types.ModuleType(name)
    m.__path__ = []
    sys.modules[name] = m
    return m


def _load(name, rel, package):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    mod.__package__ = package
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def load_reference():
    kaolin = _pkg("kaolin")
    kaolin._C = types.SimpleNamespace()
    render, ops = _pkg("kaolin.render"), _pkg("kaolin.ops")
    kaolin.render, kaolin.ops = render, ops
    _pkg("kaolin.render.mesh")
    batch = _pkg("kaolin.ops.batch")              # trianglemesh imports three names it does not use here
    batch.tile_to_packed = batch.packed_to_padded = batch.get_first_idx = None
    camera = _load("kaolin.render.camera", "kaolin/render/camera/legacy.py", "kaolin.render")
    render.camera = camera
    mesh_mod = _load("kaolin.ops.mesh.mesh", "kaolin/ops/mesh/mesh.py", "kaolin.ops.mesh")
    tri_mod = _load("kaolin.ops.mesh.trianglemesh", "kaolin/ops/mesh/trianglemesh.py", "kaolin.ops.mesh")
    opsmesh = _pkg("kaolin.ops.mesh")
    opsmesh.index_vertices_by_faces = mesh_mod.index_vertices_by_faces
    opsmesh.face_normals = tri_mod.face_normals
    ops.mesh = opsmesh
    utils = _load("kaolin.render.mesh.utils", "kaolin/render/mesh/utils.py", "kaolin.render.mesh")
    return utils, camera


def main():
    utils, camera = load_reference()
    mesh = load_obj(os.path.join(REF, "tests/samples/model.obj"))
    out = {}
    faces = torch.from_numpy(mesh["faces"]).long()
    out["faces"] = faces.numpy()
    g = torch.Generator().manual_seed(2024)
    for dname, dtype in (("f32", torch.float32), ("f64", torch.float64)):
        v = torch.from_numpy(mesh["vertices"]).to(dtype)
        v = (v - v.min(0)[0]) / (v.max() - v.min())
        B = 3
        cam_pos = torch.tensor([[0.5, 0.5, 3.], [2., 2., -2.], [3., 0.5, 0.5]], dtype=dtype)
        look_at = torch.full((B, 3), 0.5, dtype=dtype)
        up = torch.tensor([[0., 1., 0.]], dtype=dtype)
        rot, trans = camera.generate_rotate_translate_matrices(cam_pos, look_at, up)
        transform = camera.generate_transformation_matrix(cam_pos, look_at, up.repeat(B, 1))
        proj = camera.generate_perspective_projection(math.pi / 4, 1.2, dtype)
        # per-view vertices: the shared mesh, jittered per view so that the batched layout is exercised
        vb = v.unsqueeze(0).repeat(B, 1, 1) + 0.01 * torch.rand(B, v.shape[0], 3, generator=g, dtype=torch.float64).to(dtype)
        g_cam = torch.rand(B, faces.shape[0], 3, 3, generator=g, dtype=torch.float64).to(dtype)
        g_img = torch.rand(B, faces.shape[0], 3, 2, generator=g, dtype=torch.float64).to(dtype)
        g_nrm = torch.rand(B, faces.shape[0], 3, generator=g, dtype=torch.float64).to(dtype)
        for k, t in (("vertices_b", vb), ("vertices_shared", v), ("rot", rot), ("trans", trans),
                     ("transform", transform), ("proj", proj), ("g_cam", g_cam), ("g_img", g_img), ("g_nrm", g_nrm)):
            out[f"{dname}_{k}"] = t.numpy()
        for mode in ("rt", "tf"):
            for layout in ("b", "shared"):
                leaf = (vb if layout == "b" else v).clone().requires_grad_(True)
                verts = leaf if layout == "b" else leaf.unsqueeze(0).expand(B, -1, -1)
                if mode == "rt":
                    fc, fi, fn = utils.prepare_vertices(verts, faces, proj, camera_rot=rot, camera_trans=trans)
                else:
                    fc, fi, fn = utils.prepare_vertices(verts, faces, proj, camera_transform=transform)
                torch.autograd.backward([fc, fi, fn], [g_cam, g_img, g_nrm])
                key = f"{dname}_{mode}_{layout}_"
                out[key + "cam"], out[key + "img"], out[key + "nrm"] = fc.detach().numpy(), fi.detach().numpy(), fn.detach().numpy()
                out[key + "grad_vertices"] = leaf.grad.numpy()
    np.savez_compressed(os.path.join(HERE, "prepare_vertices.npz"), **out)
    print("wrote prepare_vertices.npz:", {k: v.shape for k, v in list(out.items())[:12]})


if __name__ == "__main__":
    main()
