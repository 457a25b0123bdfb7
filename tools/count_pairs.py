import torch, sys
sys.path.insert(0, '/root/repo')
import kaolin_b200 as kal
from kaolin_b200 import synth
for name, batch in (("c1", None), ("c2", 2), ("c3", 1), ("c4", None), ("c5", 1)):
    try:
        d = synth.make_inputs(name, batch=batch) if name != "c4" else synth.make_inputs("c4", views=[0, 5])
        z = d["face_vertices_z"].cuda(); img = d["face_vertices_image"].cuda()
        feats = d["face_features"][0].cuda()
        valid = None if d["face_normals_z"] is None else (d["face_normals_z"].cuda() >= 0)
        H, W = d["H"], d["W"]
        out, idx = kal.render.mesh.rasterize(H, W, z, img, feats, valid, 1000., 1e-8)
        img_m = (img * 1000.).contiguous()
        bl = d["boxlen"] * 1000.
        bb = torch.cat([img_m.min(-2)[0] - bl, img_m.max(-2)[0] + bl], -1).contiguous()
        s, p, c, t = kal._C.render.mesh.dibr_soft_mask_forward_cuda(img_m, bb, idx, d["sigmainv"], d["knum"], 1000.)
        P = idx.numel(); unc = int((idx < 0).sum()); pairs = int((c >= 0).sum())
        nz = int((p > 0).sum())
        print(name, "pixels", P, "uncovered", unc, "pairs", pairs, "pairs/pixel %.3f" % (pairs / P), "pairs/uncovered %.3f" % (pairs / max(unc, 1)), "prob>0: %.3f" % (nz / max(pairs, 1)), flush=True)
    except Exception as e:
        print(name, "ERR", e)
