"""Synthetic workloads of BASELINE.json `configs` (definitions: SURVEY.md §8(d)).

Everything is generated with torch on the CPU from ``torch.manual_seed(1000 + config_id)`` (so the GPU
path and the CPU oracle see identical tensors) and follows the reference's camera conventions:
``generate_perspective_projection(fovy=pi/4)``, look-at, ``rotate_translate_points`` and
``perspective_camera`` of kaolin/render/camera/legacy.py:22-37,120-159 and ``prepare_vertices`` of
kaolin/render/mesh/utils.py:128-175 (restated here in a few lines of torch; input producers, not hot path).
"""
import math

import torch

CONFIGS = {
    # name: (config_id, B, H, W, C, description)
    "c1": dict(cid=1, B=1, H=256, W=256, C=3, mesh="icosphere3", culling=True,
               desc="icosphere 1280 faces, batch 1, 256x256, 3-ch vertex colours"),
    "c2": dict(cid=2, B=16, H=512, W=512, C=9, mesh="soup20k", culling=False,
               desc="random triangle soup 20k faces, batch 16, 512x512, UV+normal+colour+mask (9 ch)"),
    "c3": dict(cid=3, B=64, H=1024, W=1024, C=3, mesh="uvsphere250x200", culling=True,
               desc="100k-face sphere, batch 64 views, 1024x1024, sigmainv 7000"),
    "c4": dict(cid=4, B=128, H=512, W=512, C=3, mesh="uvsphere200x100", culling=True,
               desc="40k-face mesh, 128 cameras, 512x512 (multi-view optimisation)"),
    "c5": dict(cid=5, B=8, H=2048, W=2048, C=3, mesh="uvsphere1000x500", culling=True,
               desc="stress: 1M-face mesh, batch 8, 2048x2048"),
}
SIGMAINV, BOXLEN, KNUM, MULTIPLIER, EPS = 7000., 0.02, 30, 1000., 1e-8


# ------------------------------------------------------------------ meshes
def icosphere(subdiv):
    t = (1.0 + math.sqrt(5.0)) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
         (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2),
         (10, 7, 6), (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11),
         (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    v = [tuple(c / math.sqrt(1 + t * t) for c in p) for p in v]
    for _ in range(subdiv):
        cache, nf = {}, []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = [(v[a][k] + v[b][k]) / 2 for k in range(3)]
                n = math.sqrt(sum(c * c for c in m))
                v.append(tuple(c / n for c in m))
                cache[key] = len(v) - 1
            return cache[key]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    return torch.tensor(v, dtype=torch.float32), torch.tensor(f, dtype=torch.long)


def uv_sphere(n_lon, n_lat, noise=0.05, gen=None):
    """Open lat/long grid of n_lon x n_lat quads -> 2*n_lon*n_lat triangles, radial noise 1 + noise*U(-1,1)."""
    lon = torch.arange(n_lon + 1, dtype=torch.float64) * (2 * math.pi / n_lon)
    lat = (torch.arange(n_lat + 1, dtype=torch.float64) + 0.0) * (math.pi / n_lat)
    la, lo = torch.meshgrid(lat, lon, indexing="ij")
    r = 1.0 + noise * (2 * torch.rand(la.shape, generator=gen, dtype=torch.float64) - 1)
    r[:, -1] = r[:, 0]
    r[0, :] = r[0, 0]
    r[-1, :] = r[-1, 0]
    v = torch.stack([r * torch.sin(la) * torch.cos(lo), r * torch.cos(la), r * torch.sin(la) * torch.sin(lo)], -1)
    v = v.reshape(-1, 3).float()
    i = torch.arange(n_lat).reshape(-1, 1) * (n_lon + 1) + torch.arange(n_lon).reshape(1, -1)
    a, b, c, d = i, i + 1, i + n_lon + 1, i + n_lon + 2
    f = torch.stack([torch.stack([a, c, b], -1), torch.stack([b, c, d], -1)], 2).reshape(-1, 3)
    return v, f.long()


# ------------------------------------------------------------------ cameras (reference conventions)
def look_at_rotation(eye, at, up):
    # kaolin/render/camera/legacy.py:39-83
    cz = at - eye
    cz = cz / (cz.norm(dim=1, keepdim=True) + 1e-10)
    cx = torch.cross(cz, up.expand_as(cz), dim=1)
    cx = cx / (cx.norm(dim=1, keepdim=True) + 1e-10)
    cy = torch.cross(cx, cz, dim=1)
    cy = cy / (cy.norm(dim=1, keepdim=True) + 1e-10)
    return torch.stack([cx, cy, -cz], dim=1), eye


def orbit_cameras(B, radius=3.0, dtype=torch.float32):
    i = torch.arange(B, dtype=torch.float64)
    az = 2 * math.pi * i / B
    el = math.radians(20.0) * torch.sin(4 * math.pi * i / B)
    eye = torch.stack([radius * torch.cos(el) * torch.sin(az), radius * torch.sin(el),
                       radius * torch.cos(el) * torch.cos(az)], -1).to(dtype)
    return look_at_rotation(eye, torch.zeros(B, 3, dtype=dtype), torch.tensor([[0., 1., 0.]], dtype=dtype))


def prepare_vertices(vertices, faces, rot, trans, fovy=math.pi / 4):
    """kaolin/render/mesh/utils.py:128-175 with legacy.py:22-37,120-159.  vertices (B,V,3) or (1,V,3)."""
    vcam = torch.matmul(vertices - trans.view(-1, 1, 3), rot.permute(0, 2, 1))
    tanfov = math.tan(fovy / 2.0)
    proj = torch.tensor([1.0 / tanfov, 1.0 / tanfov, -1.0], dtype=vertices.dtype, device=vertices.device)
    p = vcam * proj.view(1, 1, 3)
    vimg = p[:, :, :2] / p[:, :, 2:3]
    fcam = vcam[:, faces]                       # (B,F,3,3)
    fimg = vimg[:, faces]                       # (B,F,3,2)
    e1 = fcam[:, :, 1] - fcam[:, :, 0]
    e2 = fcam[:, :, 2] - fcam[:, :, 0]
    nrm = torch.cross(e1, e2, dim=-1)
    nrm = nrm / (nrm.norm(dim=-1, keepdim=True) + 1e-10)
    return fcam, fimg, nrm


# ------------------------------------------------------------------ configs
def make_inputs(name, batch=None, seed_offset=0, views=None):
    """Returns a dict of CPU float32 tensors: face_vertices_z (B,F,3), face_vertices_image (B,F,3,2),
    face_features (list of (B,F,3,c_i)), face_normals_z (B,F) or None, plus scalars.
    ``batch`` overrides B (cameras are still laid out for the config's nominal B unless ``views`` is given
    as the explicit list/range of view indices — used to shard a shared-mesh workload across ranks)."""
    cfg = dict(CONFIGS[name])
    B_nom = cfg["B"]
    B = B_nom if batch is None else batch
    gen = torch.Generator().manual_seed(1000 + cfg["cid"] + seed_offset)
    H, W = cfg["H"], cfg["W"]
    out = dict(cfg=cfg, H=H, W=W, sigmainv=SIGMAINV, boxlen=BOXLEN, knum=KNUM, multiplier=MULTIPLIER, eps=EPS)
    if cfg["mesh"].startswith("soup"):
        F = 20000
        ctr = torch.rand(B, F, 1, 2, generator=gen) * 1.8 - 0.9
        th0 = torch.rand(B, F, 1, generator=gen) * (2 * math.pi)
        th = th0 + torch.arange(3).view(1, 1, 3) * (2 * math.pi / 3) + (torch.rand(B, F, 3, generator=gen) - 0.5)
        r = 0.02 * (0.5 + torch.rand(B, F, 1, generator=gen))
        img = ctr + (r.unsqueeze(-1) * torch.stack([torch.cos(th), torch.sin(th)], -1))
        zc = -3.0 + 2.0 * torch.rand(B, F, 1, generator=gen)
        z = zc + (torch.rand(B, F, 3, generator=gen) * 0.04 - 0.02)
        feats = [torch.rand(B, F, 3, 2, generator=gen), torch.rand(B, F, 3, 3, generator=gen) * 2 - 1,
                 torch.rand(B, F, 3, 3, generator=gen), torch.ones(B, F, 3, 1)]
        out.update(face_vertices_z=z.contiguous(), face_vertices_image=img.contiguous(), face_features=feats,
                   face_normals_z=None)
        return out
    if cfg["mesh"] == "icosphere3":
        v, f = icosphere(3)
    else:
        n_lon, n_lat = (int(t) for t in cfg["mesh"][len("uvsphere"):].split("x"))
        v, f = uv_sphere(n_lon, n_lat, 0.05, gen)
    v = v / v.norm(dim=1).max()
    rot, eye = orbit_cameras(B_nom)
    if views is None:
        views = list(range(B)) if B <= B_nom else [i % B_nom for i in range(B)]
    views = torch.as_tensor(list(views), dtype=torch.long)
    rot, eye = rot[views], eye[views]
    fcam, fimg, nrm = prepare_vertices(v.unsqueeze(0), f, rot, eye)
    Bv = len(views)
    if cfg["C"] == 3:
        vcol = torch.rand(v.shape[0], 3, generator=gen)
        feats = [vcol[f].unsqueeze(0).expand(Bv, -1, -1, -1).contiguous()]
    else:
        raise ValueError(cfg["C"])
    out.update(face_vertices_z=fcam[..., 2].contiguous(), face_vertices_image=fimg.contiguous(),
               face_features=feats, face_normals_z=nrm[..., 2].contiguous(), vertices=v, faces=f, rot=rot, eye=eye)
    return out


def algorithmic_bytes(B, F, H, W, C, valid_faces):
    """fp32 compulsory traffic of the public API, BASELINE.md §3 / SURVEY.md §8(d)."""
    P, BF = B * H * W, B * F
    raster_fwd = BF * (36 + 12 * C) + (BF if valid_faces else 0) + P * (4 * C + 8)
    raster_bwd = P * (4 * C + 8) + 2 * BF * (24 + 12 * C)
    soft_fwd = BF * 24 + P * 12
    soft_bwd = P * 16 + 2 * BF * 24
    return dict(raster_fwd=raster_fwd, raster_bwd=raster_bwd, soft_fwd=soft_fwd, soft_bwd=soft_bwd,
                total=raster_fwd + raster_bwd + soft_fwd + soft_bwd)
