"""Generate the committed golden fixtures (run ONCE in the build container, where /root/reference exists).

    python tests/golden/make_golden.py

Writes, next to this file:

* ``scenes.npz``         — inputs of the reference's hot-path tests (tests/python/kaolin/render/mesh/
                           test_rasterization.py:36-134, test_dibr.py:43-62,211-267), produced with the
                           reference's OWN camera code (kaolin/render/camera/legacy.py, loaded by path) from
                           tests/samples/model.obj, for float32 and float64, flip in {0,1}, B=3.
* ``naive_raster.npz``   — outputs (face_idx, features, input gradients for a stored grad_out) of the
                           reference's pure-torch ``_naive_deftet_sparse_render`` (kaolin/render/mesh/deftet.py:101,
                           loaded by path with a stub ``kaolin._C``) on those scenes: what
                           test_rasterization.py:136-288 compares the CUDA op against.
* ``dibr_simple.npz`` / ``dibr_sphere.npz`` — the reference's golden ``.pt`` vectors
                           tests/samples/dibr/{simple,sphere}/*.pt (from Kaolin v0.10.0, test_dibr.py:79) re-encoded
                           compactly (values unchanged: float64 -> float64 for soft/prob/grad, integer tensors as
                           int16/uint8).

Nothing in tests/, bench.py or smoke() reads /root/reference at run time; they read these files.
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
from helpers import load_obj  # noqa: E402  (our own 30-line OBJ reader)


def _load_by_path(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def load_reference_modules():
    kaolin = types.ModuleType("kaolin")
    kaolin._C = types.SimpleNamespace()      # the naive renderer never calls _C
    sys.modules["kaolin"] = kaolin
    deftet = _load_by_path("kaolin_ref_deftet", os.path.join(REF, "kaolin/render/mesh/deftet.py"))
    legacy = _load_by_path("kaolin_ref_legacy", os.path.join(REF, "kaolin/render/camera/legacy.py"))
    return deftet, legacy


def index_vertices_by_faces(vf, faces):
    # kaolin/ops/mesh/mesh.py:24-45
    inp = vf.unsqueeze(2).expand(-1, -1, faces.shape[-1], -1)
    idx = faces[None, ..., None].expand(vf.shape[0], -1, -1, vf.shape[-1])
    return torch.gather(input=inp, index=idx, dim=1)


def build_scene(legacy, mesh, dtype, flip, batch_size=3):
    """test_rasterization.py:36-134 / test_dibr.py:211-267, on CPU."""
    faces = torch.from_numpy(mesh["faces"]).long()
    fuv = torch.from_numpy(mesh["face_uvs_idx"]).long()
    if flip:
        faces = torch.flip(faces, dims=(-1,))
        fuv = torch.flip(fuv, dims=(-1,))
    camera_pos = torch.tensor([[0.5, 0.5, 3.], [2., 2., -2.], [3., 0.5, 0.5]], dtype=dtype)[:batch_size]
    look_at = torch.full((batch_size, 3), 0.5, dtype=dtype)
    camera_up = torch.tensor([[0., 1., 0.]], dtype=dtype).repeat(batch_size, 1)
    camera_proj = legacy.generate_perspective_projection(fovyangle=math.pi / 4., dtype=dtype)
    vertices = torch.from_numpy(mesh["vertices"]).to(dtype).unsqueeze(0)
    vmin = vertices.min(dim=1, keepdims=True)[0]
    vmax = vertices.max(dim=1, keepdims=True)[0]
    vertices = (vertices - vmin) / (vmax - vmin)
    rot, trans = legacy.generate_rotate_translate_matrices(camera_pos, look_at, camera_up)
    vcam = legacy.rotate_translate_points(vertices, rot, trans)
    vimg = legacy.perspective_camera(vcam, camera_proj)
    fz = index_vertices_by_faces(vcam[:, :, -1:], faces).squeeze(-1)
    fimg = index_vertices_by_faces(vimg, faces)
    min_z = fz.reshape(batch_size, -1).min(dim=1, keepdims=True)[0]
    max_z = fz.reshape(batch_size, -1).max(dim=1, keepdims=True)[0]
    valid = torch.all(fz < ((min_z + max_z) / 2.).unsqueeze(-1), dim=-1)
    uvs = torch.from_numpy(mesh["uvs"]).to(dtype).unsqueeze(0)
    face_uvs = index_vertices_by_faces(uvs, fuv).repeat(batch_size, 1, 1, 1)
    rmin = vcam[:, :, -1].min(dim=1)[0]
    rmax = vcam[:, :, -1].max(dim=1)[0]
    render_range = torch.stack([rmin - 1e-2, rmax + 1e-2], dim=-1)
    return dict(face_vertices_z=fz.contiguous(), face_vertices_image=fimg.contiguous(),
                valid_faces=valid, face_uvs=face_uvs.contiguous(), render_range=render_range)


def pixel_coords(batch_size, height, width, dtype):
    # test_rasterization.py:118-125
    x = (2 * torch.arange(width, dtype=dtype) + 1 - width) / width
    y = (height - 2 * torch.arange(height, dtype=dtype) - 1.) / height
    return torch.stack([x.reshape(1, 1, -1).repeat(batch_size, height, 1),
                        y.reshape(1, -1, 1).repeat(batch_size, 1, width)], dim=-1).reshape(batch_size, -1, 2)


def main():
    torch.manual_seed(0)
    deftet, legacy = load_reference_modules()
    mesh = load_obj(os.path.join(REF, "tests/samples/model.obj"))
    assert mesh["vertices"].shape == (482, 3) and mesh["faces"].shape == (960, 3)
    H, W, B = 35, 31, 3

    scenes = {"mesh_vertices": mesh["vertices"], "mesh_faces": mesh["faces"].astype(np.int32),
              "mesh_uvs": mesh["uvs"], "mesh_face_uvs_idx": mesh["face_uvs_idx"].astype(np.int32)}
    naive = {}
    for dname, dtype in (("f32", torch.float32), ("f64", torch.float64)):
        for flip in (0, 1):
            sc = build_scene(legacy, mesh, dtype, bool(flip), B)
            key = f"{dname}_flip{flip}"
            for k, v in sc.items():
                scenes[f"{key}_{k}"] = v.numpy()
            pc = pixel_coords(B, H, W, dtype)
            rr = sc["render_range"].unsqueeze(1).repeat(1, H * W, 1)
            grad_out = torch.rand((B, H, W, 3), dtype=dtype)
            naive[f"{key}_grad_out"] = grad_out.numpy()
            for wv in (0, 1):
                kwargs = {"valid_faces": sc["valid_faces"]} if wv else {}
                z = sc["face_vertices_z"].clone().requires_grad_(True)
                img = sc["face_vertices_image"].clone().requires_grad_(True)
                uv = sc["face_uvs"].clone().requires_grad_(True)
                mask = torch.ones_like(uv[..., :1]).requires_grad_(True)
                (g_uv, g_mask), g_idx = deftet._naive_deftet_sparse_render(
                    pc, rr, z, img, [uv, mask], 1, **kwargs)
                feat = torch.cat([g_uv.reshape(B, H, W, 2), g_mask.reshape(B, H, W, 1)], dim=-1)
                feat.backward(grad_out)
                k2 = f"{key}_valid{wv}"
                naive[f"{k2}_face_idx"] = g_idx.reshape(B, H, W).numpy().astype(np.int16)
                naive[f"{k2}_features"] = feat.detach().numpy()
                naive[f"{k2}_grad_image"] = img.grad.numpy()
                naive[f"{k2}_grad_uvs"] = uv.grad.numpy()
                naive[f"{k2}_grad_mask"] = mask.grad.numpy()
                assert z.grad is None or bool(torch.all(z.grad == 0))
                print("naive", k2, "covered", int((g_idx >= 0).sum()), flush=True)
    np.savez_compressed(os.path.join(HERE, "scenes.npz"), **scenes)
    np.savez_compressed(os.path.join(HERE, "naive_raster.npz"), **naive)

    for sub, out in (("simple", "dibr_simple.npz"), ("sphere", "dibr_sphere.npz")):
        d = os.path.join(REF, "tests/samples/dibr", sub)
        g = {}
        for fn in sorted(os.listdir(d)):
            if not fn.endswith(".pt") or fn.startswith("new_face_idx"):
                continue
            t = torch.load(os.path.join(d, fn), map_location="cpu")
            a = t.numpy()
            name = fn[:-3]
            if name.startswith("close_face_idx"):
                assert np.all(a == np.round(a)) and a.max() < 32000
                a = a.astype(np.int16)              # stored 1-based, 0 = none (test_dibr.py:84-90)
            elif name.startswith("close_face_dist_type"):
                assert np.all(a == np.round(a)) and a.max() < 256
                a = a.astype(np.uint8)
            g[name] = a
            print(sub, name, a.shape, a.dtype)
        np.savez_compressed(os.path.join(HERE, out), **g)


if __name__ == "__main__":
    main()
