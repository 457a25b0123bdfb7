#!/usr/bin/env python
"""bench.py — rasterized Mpixels/s (fwd+bwd of rasterize + dibr_soft_mask = dibr_rasterization) on N B200s, one JSON
line.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3] [--impl ours|reference]

Default workload = the north_star Target (BASELINE.json configs[2]): 100k-face mesh, 64 views of 1024x1024,
forward + backward of rasterize + dibr_soft_mask.  The 64 views are a FIXED job sharded over the ranks with
kaolin_b200.multiview.shard_range (64 on one GPU, 8 per GPU at N=8): strong scaling, no data-path collective.

A "step" = one pass of the hot path over the rank's shard of views: forward through the public API, then backward
of both outputs with resident upstream gradients.
  value  device-timed (CUDA events per step, L2 flushed between steps), inputs resident in HBM.
  e2e    the same job through the public API from HOST buffers, copies inside the timed region, every step:
         H2D of the step's inputs from pinned memory — for the shared-mesh workloads (c1/c3/c4/c5) that is the mesh
         (vertices, faces, vertex colours) and the cameras, with prepare_vertices run on the device; for c2 (an
         independent soup per view) the face tensors — and D2H of what a renderer returns: interpolated features,
         soft mask, face_idx (+ the leaf gradients).  Three streams, double-buffered both ways.
  c4     (extra key of the default run) the multi-view optimisation workload: 128 cameras in total sharded over
         the ranks, one NCCL all-reduce of vertices.grad per step, with the all-reduce's own CUDA-event time.
Multi-GPU: one process per GPU (torchrun), barrier + synchronize on both sides of the timed region, max over ranks.
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
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c3", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--batch", type=int, default=None,
                    help="override the TOTAL number of views of the workload (sharded over the ranks)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-c4", action="store_true", help="skip the extra c4 (multi-view optimisation) result")
    ap.add_argument("--graphs", type=int, default=1,
                    help="1: replay the step from a CUDA graph captured over the public API calls; 0: eager launches")
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


def workload_name(workload, total):
    from kaolin_b200 import synth
    cfg = synth.CONFIGS[workload]
    s = f"{workload}: {cfg['desc']}"
    return s if total == cfg["B"] else s + f" [total views overridden: {total}]"


# ----------------------------------------------------------------------------------------------- CPU arm
def cpu_reference_run(workload, steps, warmup, total, emit=True):
    """The reference path on host cores: the C++ restatement of the reference kernels (oracle/, kind 'port' —
    the reference has no CPU implementation of these ops, rasterization.cpp:100-102), all host threads, on a
    bounded sample of the workload: ONE view of the batch per step (a stated row band for c5, whose single view
    is 4e12 pixel-face tests)."""
    import torch
    from kaolin_b200 import synth
    from oracle import build as obuild
    from oracle import oracle as O
    obuild.build_oracle()
    cores = len(os.sched_getaffinity(0))
    O.set_num_threads(cores)
    views = 1
    d = synth.make_inputs(workload, batch=views)
    H, W = d["H"], d["W"]
    rows = None
    if workload == "c5":
        band = 32
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
        line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": 0, "steps": steps, "warmup": warmup,
                "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "impl": "reference",
                "config": {"workload": workload_name(workload, total), "sample": sample},
                "cpu_baseline": base,
                "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "images_per_s": views * (n_rows / H) / dt, "gpu_launches": 0}
        print(json.dumps(line), flush=True)
    return base


# ----------------------------------------------------------------------------------------------- host placement
def bind_to_gpu_numa_node(dev_index):
    """Pinned host buffers are placed by first touch: allocate them while the process runs on the CPUs next to its
    GPU (NVML's ideal affinity), so that eight ranks do not stream through one socket's memory.  Returns the
    previous affinity (restored by the caller after the allocations) or None when NVML / the cpuset refuses."""
    try:
        import pynvml
        import torch
        prev = os.sched_getaffinity(0)
        pynvml.nvmlInit()
        p = torch.cuda.get_device_properties(dev_index)
        bus = f"{p.pci_domain_id:08x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        n_words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = {w * 64 + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        cpus &= prev
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return prev
    except Exception:
        return None


# ----------------------------------------------------------------------------------------------- GPU arm
KERNELS = ("raster_fwd", "raster_bwd", "soft_fwd", "soft_bwd")


class RenderWorkload:
    """c1/c2/c3/c5: forward rasterize + dibr_soft_mask through the public API and backward of both outputs with
    resident upstream gradients, on this rank's shard [lo, hi) of the workload's views."""

    def __init__(self, args, rank, world, dev):
        import torch
        from kaolin_b200 import multiview as mv
        from kaolin_b200 import synth
        self.torch = torch
        cfg = synth.CONFIGS[args.workload]
        self.total = args.batch or cfg["B"]
        lo, hi = mv.shard_range(self.total, rank, world)
        if hi <= lo:
            raise SystemExit(f"bench.py: workload {args.workload} has {self.total} views: nothing to do for rank {rank}")
        self.shared_mesh = not cfg["mesh"].startswith("soup")
        if self.shared_mesh:
            d = synth.make_inputs(args.workload, batch=hi - lo, views=[v % cfg["B"] for v in range(lo, hi)])
        else:                        # an independent soup per view: generate the whole job, keep the shard
            d = synth.make_inputs(args.workload, batch=self.total)
            for k in ("face_vertices_z", "face_vertices_image"):
                d[k] = d[k][lo:hi].contiguous()
            d["face_features"] = [f[lo:hi].contiguous() for f in d["face_features"]]
        self.d, self.dev, self.views = d, dev, (lo, hi)
        self.H, self.W = d["H"], d["W"]
        self.B, self.F = d["face_vertices_z"].shape[0], d["face_vertices_z"].shape[1]
        self.C = sum(f.shape[-1] for f in d["face_features"])
        self.culling = d["face_normals_z"] is not None
        g = torch.Generator().manual_seed(4242 + rank)
        self.g_feats = [torch.rand(self.B, self.H, self.W, f.shape[-1], generator=g).to(dev) for f in d["face_features"]]
        self.g_soft = torch.rand(self.B, self.H, self.W, generator=g).to(dev)
        self.d_z, self.d_img = d["face_vertices_z"].to(dev), d["face_vertices_image"].to(dev)
        self.d_feat = [f.contiguous().to(dev) for f in d["face_features"]]
        self.d_nz = None if not self.culling else d["face_normals_z"].to(dev)
        self.ones_nz = None if self.culling else torch.ones(self.B, self.F, device=dev)
        self.desc = {"total_views": self.total, "per_gpu_batch": self.B, "faces": self.F, "height": self.H,
                     "width": self.W, "channels": self.C, "knum": d["knum"], "sigmainv": d["sigmainv"],
                     "sharding": "the job's views are sharded contiguously over the ranks (multiview.shard_range); "
                                 "no data-path collective"}

    # ---- the hot path through the public API
    def _render(self, z, img, feat, nz, loss=False):
        import kaolin_b200 as kal
        d = self.d
        feats = feat if len(feat) > 1 else feat[0]
        # dibr_rasterization = rasterize(valid_faces = normals_z >= 0) + dibr_soft_mask (dibr.py:119-209); workloads
        # without culling pass normals_z = 1 (every face valid)
        out, soft, face_idx = kal.render.mesh.dibr_rasterization(
            self.H, self.W, z, img, feats, self.ones_nz if nz is None else nz, d["sigmainv"], d["boxlen"], d["knum"],
            d["multiplier"], d["eps"])
        outs = list(out) if isinstance(out, tuple) else [out]
        self.torch.autograd.backward(outs + [soft], self.g_feats + [self.g_soft])
        return outs, soft, face_idx

    def resident_step(self):
        img = self.d_img.detach().requires_grad_(True)
        feat = [f.detach().requires_grad_(True) for f in self.d_feat]
        self._render(self.d_z, img, feat, self.d_nz)

    # ---- end-to-end leg ------------------------------------------------------------------------------------
    def e2e_prepare(self, use_graphs):
        """Two buffer sets: pinned host inputs -> device inputs -> [public API step] -> device outputs -> pinned host
        outputs.  The step of set s is captured once (CUDA graph) on its static buffers."""
        torch, dev, d = self.torch, self.dev, self.d
        prev_aff = bind_to_gpu_numa_node(dev.index)
        pin = lambda t: t.contiguous().pin_memory()          # noqa: E731
        if self.shared_mesh:
            import math
            t = 1.0 / math.tan(math.pi / 8)
            self.proj = torch.tensor([t, t, -1.0], device=dev)
            g = torch.Generator().manual_seed(1000 + d["cfg"]["cid"])
            vcol = torch.rand(d["vertices"].shape[0], 3, generator=g)     # per-vertex colours (the C = 3 features)
            self.h_in = [pin(d["vertices"].unsqueeze(0)), pin(d["faces"]), pin(vcol), pin(d["rot"]), pin(d["eye"])]
        else:
            self.h_in = [pin(d["face_vertices_z"]), pin(d["face_vertices_image"])] + [pin(f) for f in d["face_features"]]
        self.h2d = sum(t.numel() * t.element_size() for t in self.h_in)
        self.copy_in, self.copy_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        self.stage_in = [[torch.empty_like(t, device=dev) for t in self.h_in] for _ in range(2)]
        for st in self.stage_in:
            for a, b in zip(st, self.h_in):
                a.copy_(b)
        self.e2e_fns, self.stage_out, self.h_out = [], [], []
        for s in range(2):
            fn = (lambda s=s: self._e2e_step(self.stage_in[s]))
            if use_graphs:
                from kaolin_b200.graphs import GraphedStep
                fn = GraphedStep(fn)
                outs = fn.outputs
            else:
                outs = fn()
            self.e2e_fns.append(fn)
            self.stage_out.append(list(outs))
            self.h_out.append([torch.empty(o.shape, dtype=o.dtype).pin_memory() for o in outs])
        self.d2h = sum(t.numel() * t.element_size() for t in self.h_out[0])
        self.copied = [torch.cuda.Event() for _ in range(2)]
        self.computed = [torch.cuda.Event() for _ in range(2)]
        self.returned = [torch.cuda.Event() for _ in range(2)]
        self.use_graphs = use_graphs
        if prev_aff is not None:
            os.sched_setaffinity(0, prev_aff)
        self.desc["e2e_pipeline"] = (
            ("inputs per step = mesh (vertices, faces, vertex colours) + this shard's cameras from pinned host memory, "
             "prepare_vertices + feature gather on the device; " if self.shared_mesh else
             "inputs per step = this shard's face tensors from pinned host memory; ") +
            "outputs per step copied back to pinned host memory: interpolated features, soft mask, face_idx (int64) and "
            "the leaf gradients; H2D, kernels and D2H of consecutive steps overlap on three streams (double-buffered), "
            "all inside one timed region")

    def _e2e_step(self, st):
        """One end-to-end step on device-resident copies of the host inputs; returns what goes back to the host."""
        torch = self.torch
        if self.shared_mesh:
            from kaolin_b200.render.mesh import prepare_vertices
            verts, faces, vcol, rot, eye = st
            verts = verts.detach().requires_grad_(True)
            fcam, fimg, nrm = prepare_vertices(verts, faces, self.proj, camera_rot=rot, camera_trans=eye)
            feat = [vcol[faces].unsqueeze(0).expand(self.B, -1, -1, -1).contiguous()]
            outs, soft, face_idx = self._render(fcam[..., 2], fimg, feat, nrm[..., 2])
            return (*outs, soft, face_idx, verts.grad)
        z, img = st[0], st[1].detach().requires_grad_(True)
        feat = [f.detach().requires_grad_(True) for f in st[2:]]
        outs, soft, face_idx = self._render(z, img, feat, None)
        return (*outs, soft, face_idx, img.grad, *[f.grad for f in feat])

    def e2e_run(self, steps):
        """Enqueue `steps` pipelined end-to-end steps (the caller brackets them with events)."""
        torch, dev = self.torch, self.dev
        cur, cin, cout = torch.cuda.current_stream(dev), self.copy_in, self.copy_out
        cin.wait_stream(cur)
        cout.wait_stream(cur)
        for i in range(steps):
            s = i & 1
            with torch.cuda.stream(cin):                 # H2D of step i (overlaps the kernels of step i-1)
                if i >= 2:
                    cin.wait_event(self.computed[s])     # step i-2 is done with this input set
                for a, b in zip(self.stage_in[s], self.h_in):
                    a.copy_(b, non_blocking=True)
                self.copied[s].record(cin)
            cur.wait_event(self.copied[s])
            if i >= 2:
                cur.wait_event(self.returned[s])         # step i-2's outputs have left this output set
            outs = self.e2e_fns[s]()
            if not self.use_graphs:
                self.stage_out[s] = list(outs)
            self.computed[s].record(cur)
            with torch.cuda.stream(cout):                # D2H of step i (overlaps the kernels of step i+1)
                cout.wait_event(self.computed[s])
                for h, o in zip(self.h_out[s], self.stage_out[s]):
                    h.copy_(o, non_blocking=True)
                self.returned[s].record(cout)
        cur.wait_stream(cout)                            # the region ends when the last result is on the host

    def pixels(self):
        return self.B * self.H * self.W

    def algorithmic_bytes(self):
        from kaolin_b200 import synth
        return synth.algorithmic_bytes(self.B, self.F, self.H, self.W, self.C, self.culling)


class MultiViewWorkload:
    """c4: multi-view DIB-R shape optimisation.  The mesh is replicated, the job's cameras (128) are sharded over the
    ranks; one step = prepare_vertices -> dibr_rasterization -> losses -> backward -> ONE all-reduce of vertices.grad
    -> Adam."""

    def __init__(self, args, rank, world, dev, total=None):
        import torch
        from kaolin_b200 import multiview as mv
        from kaolin_b200 import synth
        self.torch = torch
        cfg = synth.CONFIGS["c4"]
        self.total = total or cfg["B"]
        lo, hi = mv.shard_range(self.total, rank, world)
        if hi <= lo:
            raise SystemExit(f"bench.py: c4 has {self.total} cameras: nothing to do for rank {rank}")
        per_rank = hi - lo
        d = synth.make_inputs("c4", views=[v % cfg["B"] for v in range(lo, hi)])
        self.d, self.dev = d, dev
        self.H, self.W = d["H"], d["W"]
        self.B, self.F, self.C, self.culling = per_rank, d["faces"].shape[0], 3, True
        g = torch.Generator().manual_seed(77 + rank)
        vcol = torch.rand(d["vertices"].shape[0], 3, generator=g)
        self.h_rot, self.h_eye = d["rot"].contiguous().pin_memory(), d["eye"].contiguous().pin_memory()
        tm = (torch.rand(per_rank, self.H, self.W, generator=g) > 0.5).float().to(dev)
        tf = torch.rand(per_rank, self.H, self.W, 3, generator=g).to(dev)
        self.opt = mv.MultiViewOptimizer(d["vertices"].to(dev), d["faces"].to(dev), vcol.to(dev), d["rot"].to(dev),
                                         d["eye"].to(dev), tm, tf, self.H, self.W, self.total,
                                         sigmainv=d["sigmainv"], boxlen=d["boxlen"], knum=d["knum"])
        self.h_loss = torch.zeros(1).pin_memory()
        self.h_grad = torch.zeros(d["vertices"].shape).pin_memory()
        self.h2d = (self.h_rot.numel() + self.h_eye.numel()) * 4
        self.d2h = 4 + self.h_grad.numel() * 4
        self.desc = {"total_views": self.total, "per_gpu_batch": per_rank, "faces": self.F,
                     "vertices": int(d["vertices"].shape[0]), "height": self.H, "width": self.W, "channels": 3,
                     "knum": d["knum"], "sigmainv": d["sigmainv"],
                     "sharding": "the job's cameras are sharded contiguously over the ranks; one NCCL sum all-reduce of "
                                 "vertices.grad per step",
                     "step": "dibr_rasterization_from_mesh (prepare_vertices folded into K0) + mask-IoU / image-L1 "
                             "losses (fused kernels) + backward + all-reduce + fused Adam",
                     "e2e_pipeline": "optimisation loop: per step the shard's camera matrices go host->device and the "
                                     "loss + vertices.grad come back; rendered images stay on the device"}

    def resident_step(self):          # eager launches (per-kernel timing, launch counting)
        self.opt._eager_step()

    def graph_step(self):             # replay of the captured step once opt.capture() has succeeded
        self.opt.step()

    def e2e_prepare(self, use_graphs):
        pass

    def e2e_run(self, steps):
        dev = self.dev
        for _ in range(steps):      # inputs of a step = the shard's camera matrices; result = loss + leaf gradient
            self.opt.rot.copy_(self.h_rot, non_blocking=True)
            self.opt.trans.copy_(self.h_eye, non_blocking=True)
            self.h_loss.copy_(self.opt.step().reshape(1), non_blocking=True)
            self.h_grad.copy_(self.opt.last_grad, non_blocking=True)

    def pixels(self):
        return self.B * self.H * self.W

    def algorithmic_bytes(self):
        from kaolin_b200 import synth
        return synth.algorithmic_bytes(self.B, self.F, self.H, self.W, self.C, self.culling)


def gpu_run(args):
    import torch
    import torch.distributed as dist
    from kaolin_b200 import _lib, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()

    cfg = synth.CONFIGS[args.workload]
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)     # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(fn, steps, per_kernel=None):
        """K steps, each bracketed by its own event pair; L2 is flushed (512 MB memset) between steps, outside
        the pairs.  Returns total ms (max over ranks), the per-step list and the per-kernel list."""
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
        kper = [a.elapsed_time(b) for a, b in kev] if kev is not None else None
        return max_over_ranks(sum(per)), per, kper

    def timed_region(run, steps):
        """`steps` steps enqueued by run(steps), bracketed by ONE event pair and a barrier + synchronize on both
        sides; max over ranks."""
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        a.record()
        run(steps)
        b.record()
        barrier()
        return max_over_ranks(a.elapsed_time(b))

    def measure(wl, name, graphs_ok):
        """warm-up, per-kernel shares, the timed resident leg and the timed end-to-end leg of one workload."""
        warm = max(args.warmup, 3)
        for _ in range(warm):
            wl.resident_step()
        torch.cuda.synchronize()
        kshare = {}
        for k in ("setup_count", "scan", "fill") + KERNELS:
            _, _, kper = timed(wl.resident_step, 3, per_kernel=_lib.KERNEL_IDS[k])
            kshare[k] = statistics.median(kper)
        dominant = max(KERNELS, key=lambda k: kshare[k])
        n0 = _lib.kernel_launch_count()
        wl.resident_step()
        launches_per_step = _lib.kernel_launch_count() - n0
        torch.cuda.synchronize()
        resident_fn, note, use_graphs = wl.resident_step, "eager launches", False
        if args.graphs and graphs_ok:
            try:
                from kaolin_b200.graphs import GraphedStep
                resident_fn = GraphedStep(wl.resident_step)
                use_graphs = True
                note = "CUDA graph replay of the public-API step (captured once)"
            except Exception as exc:        # launch mode only — the kernels are the same either way
                torch.cuda.synchronize()
                resident_fn = wl.resident_step
                note = f"eager launches (graph capture failed: {type(exc).__name__}: {str(exc)[:120]})"
        elif args.graphs and hasattr(wl, "graph_step"):       # c4: the optimizer captures its own whole step
            wl.opt.capture()
            resident_fn, note = wl.graph_step, wl.opt.launch_mode
        elif not graphs_ok:
            note = "eager launches"
        wl.e2e_prepare(use_graphs)
        for _ in range(2):
            resident_fn()
        total_ms, per, _ = timed(resident_fn, args.steps)
        wl.e2e_run(3)
        e2e_ms = timed_region(wl.e2e_run, args.steps)
        _, _, kper = timed(wl.resident_step, min(args.steps, 5), per_kernel=_lib.KERNEL_IDS[dominant])
        return dict(ms=total_ms / args.steps, e2e_ms=e2e_ms / args.steps, kshare=kshare, dominant=dominant,
                    kms=statistics.mean(kper), launches=launches_per_step * args.steps, note=note, warm=warm)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    if args.workload == "c4":
        wl = MultiViewWorkload(args, rank, world, dev, total=args.batch)
    else:
        wl = RenderWorkload(args, rank, world, dev)
    m = measure(wl, args.workload, graphs_ok=args.workload != "c4")
    total_pixels = wl.total * wl.H * wl.W

    # ---- extra result: the multi-view optimisation workload (the only one with a collective) as a fixed 128-camera job
    c4 = None
    if args.workload == "c3" and not args.no_c4 and args.batch is None:
        wl4 = MultiViewWorkload(args, rank, world, dev)
        for _ in range(max(args.warmup, 3)):
            wl4.resident_step()
        # the collective's own time: an eager pass with CUDA events around the all-reduce (before the capture)
        wl4.opt.time_all_reduce = True
        barrier()
        for _ in range(min(args.steps, 5)):
            wl4.resident_step()
        barrier()
        ar = max_over_ranks(statistics.mean(wl4.opt.all_reduce_ms()) if world > 1 else 0.0)
        wl4.opt.time_all_reduce = False
        if args.graphs:
            wl4.opt.capture()
        for _ in range(2):
            wl4.graph_step()
        total4, _, _ = timed(wl4.graph_step, args.steps)     # per-step event pairs, L2 flushed between steps
        ms4 = total4 / args.steps
        e2e4 = timed_region(wl4.e2e_run, args.steps) / args.steps
        px4 = wl4.total * wl4.H * wl4.W
        c4 = {"workload": workload_name("c4", wl4.total), "scaling": "strong", "cameras_total": wl4.total,
              "per_gpu_cameras": wl4.B, "ms_per_step": ms4, "value": px4 / (ms4 * 1e-3) / 1e6, "unit": UNIT,
              "images_per_s": wl4.total / (ms4 * 1e-3), "all_reduce_ms": ar,
              "all_reduce": "one NCCL sum all-reduce of vertices.grad per step, CUDA events around the call on the "
                            "compute stream (max over ranks of the mean)" if world > 1 else "single rank: no collective",
              "e2e": {"ms_per_step": e2e4, "value": px4 / (e2e4 * 1e-3) / 1e6, "unit": UNIT,
                      "h2d_bytes_per_step": wl4.h2d, "d2h_bytes_per_step": wl4.d2h},
              "step": wl4.desc["step"], "launch_mode": wl4.opt.launch_mode,
              "l2": "512 MB memset between timed steps (outside the per-step event pairs)"}
    clocks = sampler.stop() if rank == 0 else None

    if rank == 0:
        ms = m["ms"]
        value = total_pixels / (ms * 1e-3) / 1e6
        e2e_value = total_pixels / (m["e2e_ms"] * 1e-3) / 1e6
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        algo = wl.algorithmic_bytes()                      # this rank's shard: what one launch processes
        dominant, kms = m["dominant"], m["kms"]
        achieved = algo[dominant] / (kms * 1e-3) / 1e9
        traffic = None
        try:       # ncu `--set full` DRAM bytes of that kernel per view, one launch (profiles/traffic.json)
            per_view = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[args.workload]
            traffic = per_view[dominant] * wl.B / per_view.get("views", cfg["B"])
        except Exception:
            pass
        config = {"workload": workload_name(args.workload, wl.total),
                  "l2": "512 MB memset between timed steps (outside the per-step event pairs)"}
        config.update(wl.desc)
        config["launch_mode"] = m["note"]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": m["warm"], "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "images_per_s": wl.total / (ms * 1e-3),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": wl.h2d, "d2h_bytes_per_step": wl.d2h,
                    "ms_per_step": m["e2e_ms"], "bytes": "per rank"},
            "gpu_launches": int(m["launches"]),
            "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "kernel_ms": kms,
                         "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650",
                         "algorithmic_bytes": algo[dominant], "views_per_launch": wl.B,
                         "whole_step": {"algorithmic_bytes": algo["total"],
                                        "achieved": algo["total"] / (ms * 1e-3) / 1e9,
                                        "frac": algo["total"] / (ms * 1e-3) / 1e9 / peak},
                         "kernel_ms_all": m["kshare"],
                         "note": "the kernels are latency / instruction-issue bound, not HBM bound (DESIGN.md §5)"},
            "clocks": clocks,
        }
        if c4 is not None:
            line["c4"] = c4
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_reference_run(args.workload, 1, 0, wl.total, emit=False)
        print(json.dumps(line), flush=True)
    if world > 1:
        # Teardown: a CUDA graph that captured an NCCL collective keeps the communicator busy, and destroying the
        # process group under it deadlocks (measured: the 2-GPU run printed its line and then hung).  Release the
        # graphs first, make sure every rank is done, and leave without running NCCL's destructors.
        if c4 is not None:
            wl4.opt._graph = None
        del wl
        import gc
        gc.collect()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def main():
    args = parse()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        from kaolin_b200 import synth
        total = args.batch or synth.CONFIGS[args.workload]["B"]
        cpu_reference_run(args.workload, max(1, min(args.steps, 3)), min(args.warmup, 1), total)
        return
    gpu_run(args)


if __name__ == "__main__":
    main()
