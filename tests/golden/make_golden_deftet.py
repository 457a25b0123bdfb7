"""Golden vectors of the DefTet sparse renderer — run ONCE in the build container, where /root/reference exists:

    python tests/golden/make_golden_deftet.py        ->  tests/golden/deftet_naive.npz

Outputs and autograd gradients of the reference's pure-torch ``_naive_deftet_sparse_render``
(kaolin/render/mesh/deftet.py:100-267), imported by path, on the reference's test mesh (tests/samples/model.obj
seen from the three cameras of tests/python/kaolin/render/mesh/test_deftet.py:326-352) at random pixel
coordinates, with knum larger than the deepest pixel (so "first knum by depth" == "first knum by mesh order",
deftet.py:112-116), full depth range and "up to the centre" range.  Nothing reads /root/reference at test time.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, os.pardir))
from make_golden import load_reference_modules, build_scene, REF  # noqa: E402
from helpers import load_obj  # noqa: E402


def main():
    torch.manual_seed(1)
    deftet, legacy = load_reference_modules()
    mesh = load_obj(os.path.join(REF, "tests/samples/model.obj"))
    B, P, KNUM = 3, 160, 24
    out = {}
    for dname, dtype in (("f32", torch.float32), ("f64", torch.float64)):
        sc = build_scene(legacy, mesh, dtype, False, B)
        pc = torch.rand((B, P, 2), dtype=torch.float64).to(dtype) * 2. - 1.
        out[f"{dname}_pixel_coords"] = pc.numpy()
        for k in ("face_vertices_z", "face_vertices_image", "face_uvs"):
            out[f"{dname}_{k}"] = sc[k].numpy()
        zmin = sc["face_vertices_z"].reshape(B, -1).min(dim=1)[0]
        zmax = sc["face_vertices_z"].reshape(B, -1).max(dim=1)[0]
        for rname, lo in (("full", zmin), ("center", (zmin + zmax) / 2.)):
            rr = torch.stack([lo, torch.zeros_like(lo)], dim=-1).unsqueeze(1).repeat(1, P, 1)
            img = sc["face_vertices_image"].clone().requires_grad_(True)
            uv = sc["face_uvs"].clone().requires_grad_(True)
            mask = torch.ones_like(uv[..., :1]).requires_grad_(True)
            (f_uv, f_mask), idx = deftet._naive_deftet_sparse_render(pc, rr, sc["face_vertices_z"], img,
                                                                     [uv, mask], KNUM)
            assert int((idx >= 0).sum(-1).max()) < KNUM, "knum must exceed the deepest pixel"
            g_uv = torch.rand(f_uv.shape, dtype=torch.float64).to(dtype)
            g_mask = torch.rand(f_mask.shape, dtype=torch.float64).to(dtype)
            torch.autograd.backward([f_uv, f_mask], [g_uv, g_mask])
            key = f"{dname}_{rname}_"
            out[key + "render_ranges"] = rr.numpy()
            out[key + "face_idx"] = idx.numpy().astype(np.int16)
            out[key + "feat_uv"], out[key + "feat_mask"] = f_uv.detach().numpy(), f_mask.detach().numpy()
            out[key + "g_uv"], out[key + "g_mask"] = g_uv.numpy(), g_mask.numpy()
            out[key + "grad_image"], out[key + "grad_uvs"] = img.grad.numpy(), uv.grad.numpy()
            out[key + "grad_mask"] = mask.grad.numpy()
            print(dname, rname, "hits", int((idx >= 0).sum()), "deepest", int((idx >= 0).sum(-1).max()), flush=True)
    np.savez_compressed(os.path.join(HERE, "deftet_naive.npz"), **out)


if __name__ == "__main__":
    main()
