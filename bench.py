#!/usr/bin/env python
"""bench.py — rasterized Mpixels/s (fwd+bwd of rasterize + dibr_soft_mask) on N B200s, one JSON line.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2] [--impl ours|reference]

A "step" = one pass of the hot path over one batch of synthetic input (BASELINE.json configs):
forward rasterize + dibr_soft_mask through the public API, then backward of both outputs with resident
upstream gradients.  `value` is device-timed (CUDA events) with inputs resident in HBM; `e2e` repeats the
measurement with the step's inputs copied from pinned host memory and a scalar result read back, inside
the timed region.  Multi-GPU: one process per GPU (torchrun), the batch of views is sharded with no data-path
collective (weak scaling: every rank renders its own batch), barrier-bracketed, max over ranks.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "rasterized Mpixels/s (rasterize + dibr_soft_mask, fwd+bwd)"
UNIT = "Mpixels/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--batch", type=int, default=None, help="override the per-rank batch of the workload")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------------------------- CPU arm
def cpu_reference_run(workload, steps, warmup, emit=True, budget_views=None):
    """The reference path on host cores: the C++ restatement of the reference kernels (oracle/, kind 'port' —
    the reference has no CPU implementation of these ops, rasterization.cpp:100-102), all host threads,
    on a bounded sample of the workload: `views` views of the batch per step."""
    import torch
    from kaolin_b200 import synth
    from oracle import build as obuild
    from oracle import oracle as O
    obuild.build_oracle()
    cores = len(os.sched_getaffinity(0))
    O.set_num_threads(cores)
    views = budget_views or {"c1": 1, "c2": 1, "c3": 1, "c4": 1, "c5": 1}[workload]
    d = synth.make_inputs(workload, batch=views)
    H, W = d["H"], d["W"]
    rows = None
    if workload in ("c3", "c5"):            # one view exceeds the time budget: a stated band of rows
        band = {"c3": 128, "c5": 32}[workload]
        rows = (H // 2 - band // 2, H // 2 + band // 2)
    z, img = d["face_vertices_z"], d["face_vertices_image"]
    feats = torch.cat(d["face_features"], -1).contiguous()
    valid = None if d["face_normals_z"] is None else d["face_normals_z"] >= 0
    n_rows = H if rows is None else rows[1] - rows[0]
    g = torch.Generator().manual_seed(1)
    g_feat = torch.rand(views, H, W, feats.shape[-1], generator=g)
    g_soft = torch.rand(views, H, W, generator=g)

    def step():
        f, idx, w = O.rasterize_forward(H, W, z, img, feats, valid, d["multiplier"], d["eps"], rows=rows)
        soft, img_m, prob, cidx, ctype = O.dibr_soft_mask_forward(img, idx, d["sigmainv"], d["boxlen"], d["knum"],
                                                                 d["multiplier"], rows=rows)
        O.rasterize_backward(g_feat, idx, w, img, feats, d["eps"], rows=rows)
        O.dibr_soft_mask_backward(g_soft, soft, img_m, idx, prob, cidx, ctype, d["sigmainv"], d["multiplier"],
                                  rows=rows)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    pixels = views * n_rows * W
    val = pixels / dt / 1e6
    sample = f"{views} view(s) of {workload}" + (f", image rows {rows[0]}..{rows[1] - 1} of {H}" if rows else "") + \
             f"; {steps} timed step(s), {dt:.3f} s each"
    base = {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
    if emit:
        cfg = d["cfg"]
        line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": 0, "steps": steps, "warmup": warmup,
                "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "impl": "reference",
                "config": {"workload": f"{workload}: {cfg['desc']}", "sample": sample},
                "cpu_baseline": base,
                "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "images_per_s": views * (n_rows / H) / dt, "gpu_launches": 0}
        print(json.dumps(line), flush=True)
    return base


# ----------------------------------------------------------------------------------------------- GPU arm
KERNEL_ALGO = {  # algorithmic bytes of each kernel = the BASELINE.md §3 term it implements
    "raster_fwd": "raster_fwd", "raster_bwd": "raster_bwd", "soft_fwd": "soft_fwd", "soft_bwd": "soft_bwd"}


def gpu_run(args):
    import torch
    import torch.distributed as dist
    import kaolin_b200 as kal
    from kaolin_b200 import _lib, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()

    d = synth.make_inputs(args.workload, batch=args.batch, seed_offset=rank * 100)
    cfg = d["cfg"]
    H, W = d["H"], d["W"]
    # pinned host copies (the e2e leg copies these every step) and resident device copies (the `value` leg)
    h_z = d["face_vertices_z"].pin_memory()
    h_img = d["face_vertices_image"].pin_memory()
    h_feat = [f.contiguous().pin_memory() for f in d["face_features"]]
    h_nz = None if d["face_normals_z"] is None else d["face_normals_z"].pin_memory()
    B, F, C = h_z.shape[0], h_z.shape[1], sum(f.shape[-1] for f in h_feat)
    P = B * H * W
    splits = [f.shape[-1] for f in d["face_features"]]
    g = torch.Generator().manual_seed(4242 + rank)
    g_feats = [torch.rand(B, H, W, c, generator=g).to(dev) for c in splits]
    g_soft = torch.rand(B, H, W, generator=g).to(dev)
    d_z, d_img, d_feat = h_z.to(dev), h_img.to(dev), [f.to(dev) for f in h_feat]
    d_nz = None if h_nz is None else h_nz.to(dev)
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)     # > 126 MB L2

    def step(z, img, feat, nz, loss=False):
        img = img.detach().requires_grad_(True)
        if isinstance(feat, list):
            feats = [f.detach().requires_grad_(True) for f in feat]
            feats = feats if len(feats) > 1 else feats[0]
        else:
            feats = feat.detach().requires_grad_(True)
        valid = None if nz is None else (nz >= 0.)
        out, face_idx = kal.render.mesh.rasterize(H, W, z, img, feats, valid, d["multiplier"], d["eps"])
        soft = kal.render.mesh.dibr_soft_mask(img, face_idx, d["sigmainv"], d["boxlen"], d["knum"], d["multiplier"])
        outs = list(out) if isinstance(out, tuple) else [out]
        torch.autograd.backward(outs + [soft], g_feats + [g_soft])
        if loss:
            fl = feats if isinstance(feats, list) else [feats]
            return (img.grad.sum() + sum(f.grad.sum() for f in fl) + soft.sum()).reshape(1)
        return None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, per_kernel=None):
        """K steps, each bracketed by its own event pair; L2 is flushed (512 MB memset) between steps, outside
        the pairs.  Returns total ms (max over ranks) and the per-step list."""
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        kev = None
        if per_kernel is not None:
            kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            for a, b in kev:
                a.record(); b.record()
        barrier()
        for i in range(steps):
            flush.zero_()
            if kev is not None:
                _lib.profile_next_launch(per_kernel, kev[i][0], kev[i][1])
            ev[i][0].record()
            fn()
            ev[i][1].record()
        barrier()
        _lib.profile_next_launch(-1, None, None)
        per = [a.elapsed_time(b) for a, b in ev]
        tot = torch.tensor([sum(per)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        kper = [a.elapsed_time(b) for a, b in kev] if kev is not None else None
        return float(tot.item()), per, kper

    resident = lambda: step(d_z, d_img, d_feat, d_nz)                                       # noqa: E731
    h_loss = torch.zeros(1).pin_memory()

    def e2e_step():
        z, img = h_z.to(dev, non_blocking=True), h_img.to(dev, non_blocking=True)
        feat = [f.to(dev, non_blocking=True) for f in h_feat]
        nz = None if h_nz is None else h_nz.to(dev, non_blocking=True)
        h_loss.copy_(step(z, img, feat, nz, loss=True), non_blocking=True)
    h2d = h_z.numel() * 4 + h_img.numel() * 4 + sum(f.numel() for f in h_feat) * 4 + \
        (0 if h_nz is None else h_nz.numel() * 4)

    # ---- warm-up (also: which kernel dominates?)
    for _ in range(max(args.warmup, 3)):
        resident()
    torch.cuda.synchronize()
    kshare = {}
    for name in ("setup_count", "scan", "fill", "raster_fwd", "raster_bwd", "soft_fwd", "soft_bwd"):
        _, _, kper = timed(resident, 3, per_kernel=_lib.KERNEL_IDS[name])
        kshare[name] = statistics.median(kper)
    dominant = max(KERNEL_ALGO, key=lambda k: kshare[k])

    # ---- the timed region
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    n0 = _lib.kernel_launch_count()
    total_ms, per, kper = timed(resident, args.steps, per_kernel=_lib.KERNEL_IDS[dominant])
    launches = _lib.kernel_launch_count() - n0
    for _ in range(2):
        e2e_step()
    e2e_ms, _, _ = timed(e2e_step, args.steps)
    clocks = sampler.stop() if rank == 0 else None

    if rank == 0:
        ms = total_ms / args.steps
        value = world * P / (ms * 1e-3) / 1e6
        e2e_value = world * P / (e2e_ms / args.steps * 1e-3) / 1e6
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        algo = synth.algorithmic_bytes(B, F, H, W, C, h_nz is not None)
        kms = statistics.mean(kper)
        achieved = algo[KERNEL_ALGO[dominant]] / (kms * 1e-3) / 1e9
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload, {}).get(dominant)
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {cfg['desc']}", "per_gpu_batch": B, "faces": F, "height": H,
                       "width": W, "channels": C, "knum": d["knum"], "sigmainv": d["sigmainv"],
                       "l2": "512 MB memset between timed steps (outside the per-step event pairs)",
                       "sharding": "views across ranks, no data-path collective"},
            "images_per_s": world * B / (ms * 1e-3),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "ms_per_step": e2e_ms / args.steps},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "kernel_ms": kms,
                         "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650",
                         "algorithmic_bytes": algo[KERNEL_ALGO[dominant]],
                         "whole_step": {"algorithmic_bytes": algo["total"],
                                        "achieved": algo["total"] / (ms * 1e-3) / 1e9,
                                        "frac": algo["total"] / (ms * 1e-3) / 1e9 / peak},
                         "kernel_ms_all": kshare},
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_reference_run(args.workload, 1, 1, emit=False)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        steps = max(1, min(args.steps, 3))
        cpu_reference_run(args.workload, steps, min(args.warmup, 1))
        return
    gpu_run(args)


if __name__ == "__main__":
    main()
