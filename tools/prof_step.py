#!/usr/bin/env python
"""One profiled step of the hot path (for ncu): warm-up steps, then ONE step between cudaProfilerStart/Stop.

  python tools/prof_step.py [--workload c2] [--batch N] [--warmup 2]
  ncu --profile-from-start off --set full --clock-control none --import-source on -o gpurun_out/prof \
      python tools/prof_step.py --workload c2

Without ncu it just runs (exit 0 = the command is safe to profile) and prints the step time.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--fused", type=int, default=1, help="1: dibr_rasterization; 0: rasterize + dibr_soft_mask")
    args = ap.parse_args()
    import torch
    import kaolin_b200 as kal
    from kaolin_b200 import synth
    dev = torch.device("cuda", 0)
    d = synth.make_inputs(args.workload, batch=args.batch)
    H, W = d["H"], d["W"]
    z = d["face_vertices_z"].to(dev)
    img0 = d["face_vertices_image"].to(dev)
    feats0 = [f.contiguous().to(dev) for f in d["face_features"]]
    nz = None if d["face_normals_z"] is None else d["face_normals_z"].to(dev)
    B = z.shape[0]
    g = torch.Generator().manual_seed(4242)
    g_feats = [torch.rand(B, H, W, f.shape[-1], generator=g).to(dev) for f in feats0]
    g_soft = torch.rand(B, H, W, generator=g).to(dev)
    ones_nz = torch.ones(z.shape[0], z.shape[1], device=dev)

    def step():
        img = img0.detach().requires_grad_(True)
        feats = [f.detach().requires_grad_(True) for f in feats0]
        feats = feats if len(feats) > 1 else feats[0]
        if args.fused:          # the call bench.py times
            out, soft, face_idx = kal.render.mesh.dibr_rasterization(
                H, W, z, img, feats, ones_nz if nz is None else nz, d["sigmainv"], d["boxlen"], d["knum"],
                d["multiplier"], d["eps"])
        else:
            valid = None if nz is None else (nz >= 0.)
            out, face_idx = kal.render.mesh.rasterize(H, W, z, img, feats, valid, d["multiplier"], d["eps"])
            soft = kal.render.mesh.dibr_soft_mask(img, face_idx, d["sigmainv"], d["boxlen"], d["knum"],
                                                  d["multiplier"])
        outs = list(out) if isinstance(out, tuple) else [out]
        torch.autograd.backward(outs + [soft], g_feats + [g_soft])

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()
    a.record()
    step()
    b.record()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print(f"step {a.elapsed_time(b):.3f} ms  ({args.workload}, B={B})")


if __name__ == "__main__":
    main()
