#!/usr/bin/env python
"""bench.py -- the headline benchmark of BASELINE.json: rasterizer forward+backward at 1080p on the 3M-Gaussian,
SH-degree-3 scene (config C3), reported as views/s over all GPUs and ms per view.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

A STEP is one batch of `--views-per-step` (default 8) distinct cameras per GPU: forward + backward of every view
through the public operator API (GaussianRasterizer + autograd), per-Gaussian gradients accumulated on the device,
and -- for N > 1 -- one sum exchange of the 59-float/Gaussian gradient bucket (view-parallel data parallelism: libgsr's own
one-launch kernel over NVLink peer memory, gsr_allreduce_sum; NCCL only bootstraps the symmetric memory). `--scaling weak`
(default) keeps the per-GPU work fixed; `--scaling strong` keeps the batch fixed at 64 views and splits it over the ranks.
Inputs (708 MB of Gaussian parameters per replica) exceed the 126 MB L2, so no explicit flush is needed between iterations.

Other workloads (`--workload`): train (config 3: + fused L1/SSIM loss per view + Adam per step), c1 (config 1: 10k Gaussians
SH0, 256x256), c2 (config 2: 500k SH3, 1080p), slam (config 5: 640x480, map growing 200k -> 400k, distCUDA2 on every
keyframe's new points), knn (distCUDA2 over 3M points). Each has the same `--impl reference` arm (the unmodified reference
CUDA build, oracle/_ref). The default line also carries, outside the headline region, `drop_in_call_shape`,
`train_step`, `ssim_ms` / `adam_ms` / `densify_ms` and -- for N > 1 -- `parity_multi_gpu` (rank 0 renders ALL views of the
step alone and compares with the exchanged bucket).

One JSON line is printed by rank 0. Keys beyond the base contract:
  roofline      dominant kernel of the step (by device time: CUDA events on the launching stream around every launch, in
                a second pass over the same K steps right after the headline region) against the ceiling that bounds it; roofline_hbm: the same for the most expensive HBM-bound
                kernel against the measured copy bandwidth of MEASURED_PEAKS.json; kernels: the full per-kernel table.
  cpu_baseline  the C restatement of the reference algorithm (oracle/, OpenMP) timed on the host cores on a bounded
                sample of the same workload (rank 0, N=1 only). Checker code: measured here, never shipped.
  e2e           the same metric through the same API with HOST buffers: every view's camera and dL/dpixel come from
                pinned host memory and the rendered image goes back to pinned host memory inside the timed region.
`--impl reference` times the UNMODIFIED reference CUDA rasterizer (oracle/_ref/ref_dgr_C.so, built from
/root/reference by oracle/Makefile) on the same step; if that build is absent it times the CPU restatement instead.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "gaussian-splatting-slam_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "rasterizer fwd+bwd views/s @1080p, 3M Gaussians SH3 (ms/iter = 1000 * n_gpus / value)"
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12      # no SIMT figure in MEASURED_PEAKS.json: 148 SM x 128 FMA/clk x max clock


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--views-per-step", type=int, default=8)
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--workload", default="raster", choices=["raster", "train", "c1", "c2", "slam", "knn"],
                    help="raster: rasterizer forward+backward (BASELINE metric); train: + fused L1/SSIM loss per view and one "
                         "Adam step per step (BASELINE config 3, 'full train-step'); c1 / c2 / slam: BASELINE configs 1, 2, 5 "
                         "with the reference's exact call shape; knn: distCUDA2 over --gaussians points")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --views-per-step views per GPU; strong: 8 x --views-per-step views in total, split over the ranks")
    ap.add_argument("--no-extras", action="store_true", help="skip the secondary measurements of the default line")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the all-views-on-rank-0 gradient check")
    ap.add_argument("--aa", action="store_true", help="c1 / c2 / slam, ours only: antialiasing + inverse-depth loss on")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-kernel-timing", action="store_true")
    ap.add_argument("--no-fused-exchange", action="store_true",
                    help="train workload on several GPUs: all-reduce + Adam on every rank instead of the one-launch exchange+Adam")
    ap.add_argument("--no-grad-sinks", action="store_true", help="let autograd accumulate (reference semantics) instead of in-kernel accumulation")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock / power / throttle reasons sampled every 100 ms WHILE the timed region runs. NVML in a thread (set up
    before the warm-up, so nothing is forked or initialised inside the region); nvidia-smi -lms as the fallback."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.nv, self.h = index, [], None, None, None
        self.stop_flag = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].strip().isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nv = pynvml
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def start(self):
        self.stop_flag.clear()
        if os.environ.get("GSR_BENCH_NO_SAMPLER"):          # diagnosis only: a line without clocks is not a bench value
            self.nv, self.proc = None, None
            return
        if self.nv is not None:
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nv
        R = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
             "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self.stop_flag.is_set():
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.rows.append([str(sm), str(self.sm_max), str(pw)] + ["Active" if bits & R[k] else "Not Active" for k in
                                 ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")])
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        self.stop_flag.set()
        if self.nv is not None:
            self.t.join(timeout=1.0)
        elif self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons),
                "source": "nvml" if self.nv is not None else "nvidia-smi"}


# ----------------------------------------------------------------------------------------------- workload
def total_views(args, world):
    """Views of one step over all ranks: weak scaling fixes the per-GPU batch, strong scaling the global one (64)."""
    return args.views_per_step * (8 if args.scaling == "strong" else world)


def build_workload(args, rank, world, dev):
    import torch
    import scenes
    sc = scenes.make_c3(P=args.gaussians)
    n_views = total_views(args, world)
    cams = scenes.c3_cameras(n=n_views, W=args.width, H=args.height, jitter=0.05 if n_views > 8 else 0.0)
    from view_parallel import shard_views
    mine = [cams[i] for i in shard_views(n_views, rank, world)]
    dpix_host = rank_dpix(args, rank, pinned=dev.type == "cuda")
    return sc, mine, dpix_host, cams


def rank_dpix(args, rank, pinned=False):
    """dL/dpixel of every view a rank renders (seeded by the rank, so that another rank can restate it)."""
    import torch
    g = torch.Generator().manual_seed(1000 + rank)
    t = torch.randn(3, args.height, args.width, generator=g)
    return t.pin_memory() if pinned else t


def device_params(sc, dev):
    import torch
    names = ("means3D", "shs", "opacities", "scales", "rotations")
    return {n: getattr(sc, n).to(dev).contiguous().requires_grad_(True) for n in names}


# The train-step workload runs the optimiser with the reference's groups (gaussian_model.py:160-167; shs = f_dc and
# f_rest together here) but learning rate 0: Adam reads and writes every parameter and both moments exactly as in
# training (same kernel, same bytes), while the synthetic scene -- and with it num_rendered and every list length --
# stays the same from the first warm-up step to the last timed one. (With the real rates and a random target image
# the scene drifts and consecutive steps time different workloads.)
TRAIN_LRS = {"means3D": 0.0, "shs": 0.0, "opacities": 0.0, "scales": 0.0, "rotations": 0.0}
# ... and ONE checked step with the reference's real rates (arguments/__init__.py:75-85: position_lr_init 1.6e-4,
# feature_lr 2.5e-3, opacity_lr 5e-2 [upstream 0.025 / 0.05 depending on the version], scaling_lr 5e-3, rotation_lr 1e-3)
# after the timed region: the parameters it produces are compared with torch.optim.Adam on the same gradients.
CHECK_LRS = {"means3D": 1.6e-4, "shs": 2.5e-3, "opacities": 5e-2, "scales": 5e-3, "rotations": 1e-3}


class OursArm:
    name = "ours"

    def __init__(self, args, sc, cams, dpix_host, dev):
        import torch
        from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, _C
        from view_parallel import GradBucket
        self.torch, self._C = torch, _C
        self.Settings, self.Rasterizer = GaussianRasterizationSettings, GaussianRasterizer
        self.dev, self.args, self.deg = dev, args, sc.sh_degree
        self.params = device_params(sc, dev)
        self.means2D = torch.zeros_like(self.params["means3D"], requires_grad=True)    # the caller's grad sink
        self.bucket = GradBucket(self.params)
        self.sinks_first = self.bucket.views_first
        self.bg = torch.zeros(3, device=dev)
        self.cams_host = cams
        self.cams_dev = [(c.viewmatrix.to(dev), c.projmatrix.to(dev), c.campos.to(dev)) for c in cams]
        self.dpix_host = dpix_host
        self.dpix = dpix_host.to(dev)
        # pinned staging for the end-to-end leg
        self.cam_pinned = [torch.cat([c.viewmatrix.reshape(-1), c.projmatrix.reshape(-1), c.campos.reshape(-1)]).pin_memory()
                           for c in cams]
        self.color_host = torch.empty(3, args.height, args.width).pin_memory()
        self.last = None
        self.train = args.workload == "train"
        self.fused_opt = False
        if self.train:
            from fused_ssim import fused_l1_ssim_loss
            import gsr_optim
            self.loss_fn = fused_l1_ssim_loss
            self.gt = torch.rand(3, args.height, args.width, generator=torch.Generator().manual_seed(7)).to(dev)
            self.fused_opt = self.bucket.peer is not None and not args.no_fused_exchange
            if self.fused_opt:
                # several GPUs: parameters move into peer memory as well and the exchange + Adam become ONE launch per
                # rank (each rank updates 1 / world of the parameters and keeps 1 / world of the moments)
                from view_parallel import ExchangeAdam, ParamBucket
                pb = ParamBucket(self.params, self.bucket)
                self.params = pb.tensors
                self.bucket.params = pb.tensors
                self.opt = ExchangeAdam(self.bucket, pb, TRAIN_LRS, eps=1e-15)
            else:
                self.opt = gsr_optim.FusedAdam([{"params": [self.params[k]], "lr": lr, "name": k} for k, lr in TRAIN_LRS.items()],
                                               lr=0.0, eps=1e-15)

    def _settings(self, cam, view, proj, campos):
        return self.Settings(image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=self.bg,
                             scale_modifier=1.0, viewmatrix=view, projmatrix=proj, sh_degree=self.deg, campos=campos,
                             prefiltered=False, debug=False, antialiasing=False)

    def _sinks(self, first):
        """grad_sinks of a view: the first view of a step overwrites the bucket (no bucket.zero()), the others add."""
        if self.args.no_grad_sinks:
            return None
        return self.sinks_first if first else self.bucket.views

    def _view(self, st, dpix, first=False):
        p = self.params
        color, radii, _invd = self.Rasterizer(st)(means3D=p["means3D"], means2D=self.means2D, opacities=p["opacities"],
                                                  shs=p["shs"], scales=p["scales"], rotations=p["rotations"],
                                                  grad_sinks=self._sinks(first))
        if self.train:
            self.loss_fn(color, self.gt, 0.2).backward()          # train.py:114-122 (lambda_dssim = 0.2)
        else:
            self.torch.autograd.backward([color], [dpix])
        self.last = (color, radii)
        return color

    def step(self):
        if self.args.no_grad_sinks:
            self.bucket.zero()                    # autograd adds into .grad: the bucket has to start at zero
        self.bucket.attach()
        self.means2D.grad = None
        for i, (cam, (v, pj, cp)) in enumerate(zip(self.cams_host, self.cams_dev)):
            self._view(self._settings(cam, v, pj, cp), self.dpix, first=(i == 0))
        if self.train and self.fused_opt:
            self.opt.step()                       # gsr_exchange_adam_step: the exchange and the optimiser in one launch
        else:
            self.bucket.all_reduce()
            if self.train:
                self.opt.step()

    def step_e2e(self):
        """Same step, host buffers: every view's camera + dL/dpixel come from pinned host memory and its rendered image goes
        back to pinned host memory. The copies ride two side streams and double buffers, so view i+1's upload and view
        i-1's download overlap view i's kernels; the step ends when all three streams have drained."""
        torch = self.torch
        if not hasattr(self, "_e2e"):
            n = 2
            self._e2e = dict(
                h2d=torch.cuda.Stream(self.dev), d2h=torch.cuda.Stream(self.dev),
                cam=[torch.empty(35, device=self.dev) for _ in range(n)],
                dpix=[torch.empty_like(self.dpix) for _ in range(n)],
                host=[torch.empty(3, self.args.height, self.args.width).pin_memory() for _ in range(n)],
                up=[torch.cuda.Event() for _ in range(n)], used=[torch.cuda.Event() for _ in range(n)],
                down=[torch.cuda.Event() for _ in range(n)])
        e = self._e2e
        main = torch.cuda.current_stream(self.dev)
        if self.args.no_grad_sinks:
            self.bucket.zero()
        self.bucket.attach()
        self.means2D.grad = None
        n_views = len(self.cams_host)

        def upload(i):
            k = i % 2
            with torch.cuda.stream(e["h2d"]):
                if i >= 2:
                    e["h2d"].wait_event(e["used"][k])            # view i-2 is done with this buffer pair
                e["cam"][k].copy_(self.cam_pinned[i], non_blocking=True)
                e["dpix"][k].copy_(self.dpix_host, non_blocking=True)
                e["up"][k].record(e["h2d"])

        upload(0)
        for i, cam in enumerate(self.cams_host):
            k = i % 2
            if i + 1 < n_views:
                upload(i + 1)
            main.wait_event(e["up"][k])
            b = e["cam"][k]
            p = self.params
            color, radii, _invd = self.Rasterizer(self._settings(cam, b[0:16].view(4, 4), b[16:32].view(4, 4), b[32:35]))(
                means3D=p["means3D"], means2D=self.means2D, opacities=p["opacities"], shs=p["shs"], scales=p["scales"],
                rotations=p["rotations"], grad_sinks=self._sinks(i == 0))
            rendered = torch.cuda.Event()
            rendered.record(main)
            with torch.cuda.stream(e["d2h"]):                     # the image goes home while the backward runs
                e["d2h"].wait_event(rendered)
                if i >= 2:
                    e["d2h"].wait_event(e["down"][k])
                e["host"][k].copy_(color.detach(), non_blocking=True)
                e["down"][k].record(e["d2h"])
                color.record_stream(e["d2h"])
            if self.train:
                self.loss_fn(color, self.gt, 0.2).backward()
            else:
                torch.autograd.backward([color], [e["dpix"][k]])
            e["used"][k].record(main)
            self.last = (color, radii)
        main.wait_stream(e["d2h"])
        if self.train and self.fused_opt:
            self.opt.step()                       # gsr_exchange_adam_step: the exchange and the optimiser in one launch
        else:
            self.bucket.all_reduce()
            if self.train:
                self.opt.step()

    def state_of_view(self, i):
        """Forward of view i through the binding (outside any timed region) and its decoded scratch state."""
        cam, (v, pj, cp) = self.cams_host[i], self.cams_dev[i]
        p = {k: t.detach() for k, t in self.params.items()}
        e = self.torch.empty(0, device=self.dev)
        R, _color, radii, geom, binning, img, _ = self._C.rasterize_gaussians(
            self.bg, p["means3D"], e, p["opacities"], p["scales"], p["rotations"], 1.0, e, v, pj, cam.tanfovx, cam.tanfovy,
            cam.H, cam.W, p["shs"], self.deg, cp, False, False, False)
        st = self._C.decode_state(self.args.gaussians, R, cam.W, cam.H, geom, binning, img)
        st["pair_counts"] = self._C.blend_stats(self.args.gaussians, R, cam.W, cam.H, geom, binning, img)
        return R, radii, st

    def e2e_bytes(self):
        n = len(self.cams_host)
        return n * (35 * 4 + self.dpix_host.numel() * 4), n * self.color_host.numel() * 4

    def launches(self):
        return self._C.launch_count()


class ReferenceArm:
    """The unmodified reference CUDA rasterizer behind the same step (autograd wrapper = harness, oracle/ref_cuda.py)."""
    name = "reference"

    def __init__(self, args, sc, cams, dpix_host, dev):
        import torch
        from oracle import ref_cuda
        from view_parallel import GradBucket
        self.torch, self.ref = torch, ref_cuda
        self.dev, self.deg = dev, sc.sh_degree
        self.params = device_params(sc, dev)
        self.means2D = torch.zeros_like(self.params["means3D"], requires_grad=True)
        self.bucket = GradBucket(self.params)
        self.bg = torch.zeros(3, device=dev)
        self.cams_host = cams
        self.cams_dev = [(c.viewmatrix.to(dev), c.projmatrix.to(dev), c.campos.to(dev)) for c in cams]
        self.dpix = dpix_host.to(dev)
        self.fn = ref_cuda.autograd_function()
        self.train = args.workload == "train"
        if self.train:
            self.ssim = ref_cuda.fused_ssim_function()               # the reference's fused-ssim CUDA build
            self.gt = torch.rand(3, args.height, args.width, generator=torch.Generator().manual_seed(7)).to(dev)
            self.opt = torch.optim.Adam([{"params": [self.params[k]], "lr": lr, "name": k} for k, lr in TRAIN_LRS.items()],
                                        lr=0.0, eps=1e-15)          # scene/gaussian_model.py:168

    def step(self):
        self.bucket.zero()
        self.bucket.attach()
        self.means2D.grad = None
        p = self.params
        for cam, (v, pj, cp) in zip(self.cams_host, self.cams_dev):
            color, _radii = self.fn.apply(p["means3D"], self.means2D, p["shs"], p["opacities"], p["scales"], p["rotations"],
                                          dict(bg=self.bg, viewmatrix=v, projmatrix=pj, campos=cp, tanfovx=cam.tanfovx,
                                               tanfovy=cam.tanfovy, H=cam.H, W=cam.W, sh_degree=self.deg))
            if self.train:                                           # train.py:114-122
                l1 = self.torch.abs(color - self.gt).mean()
                loss = 0.8 * l1 + 0.2 * (1.0 - self.ssim(color.unsqueeze(0), self.gt.unsqueeze(0)))
                loss.backward()
            else:
                self.torch.autograd.backward([color], [self.dpix])
        if self.train:
            self.opt.step()

    def launches(self):
        return 0


# ----------------------------------------------------------------------------------------------- work model
def work_model(arm, args):
    """Algorithmic work per VIEW for every kernel name (SURVEY.md 8d / DESIGN.md 'Roofline accounting'), taken from the
    state of the last rendered view. HBM kernels: bytes; blend kernels: FP32 flop."""
    import torch
    P = args.gaussians
    R, radii, st = arm.state_of_view(len(arm.cams_host) - 1)
    visible = int((radii > 0).sum())
    traversed, contributing = st["pair_counts"]          # what renderCUDA iterates / what passes its alpha test
    assert traversed == int(st["n_contrib"].to(torch.int64).sum())
    tiles = ((args.width + 15) // 16) * ((args.height + 15) // 16)
    N = args.width * args.height
    sort_tiles_P, sort_tiles_R = (P + 4095) // 4096, (R + 4095) // 4096
    tb = max(1, math.ceil(math.log2(tiles)))
    rp = (tb + 7) // 8
    w = {
        # per visible Gaussian 311 B (read 236 + write 75), per culled one 20 B
        "preprocess_fwd": ("hbm", 311.0 * visible + 20.0 * (P - visible)),
        # (P-sort: 4 passes) + (R-sort: rp passes): one histogram read of the keys per sort, then 8 B read + 8 B written
        # per pair and pass
        # only the depth sort has a histogram kernel (one read of the P depth keys); the instance sort's digit histograms
        # come from the tile table
        "radix_hist": ("hbm", 4.0 * P),
        # depth sort in drop mode: pass 1 reads all P pairs and writes the visible ones, passes 2-4 move visible pairs only
        "onesweep_depth": ("hbm", 8.0 * P + 8.0 * visible + 3 * 16.0 * visible),
        # instance sort by tile id: 8 B read + 8 B written per pair and pass; the last pass writes no keys
        "onesweep_tile": ("hbm", 16.0 * rp * R - 4.0 * R),
        # pass 2 of the two-launch scan (or the chained one-launch scan of small clouds): parked rectangle read, offset
        # written per visible Gaussian
        "scan_offsets": ("hbm", 8.0 * visible),
        # emission balanced by output: (tile id, Gaussian id) written per instance; offset, id and rectangle read per
        # visible Gaussian
        "emit_instances": ("hbm", 12.0 * visible + 8.0 * R),
        # two-launch scan, pass 1: id + gathered packed rectangle read, rectangle parked in depth order
        "scan_tile_sums": ("hbm", 12.0 * visible),
        "count_instances": ("hbm", 4.0 * P),
        # packed rectangles of all P Gaussians read once; difference array -> ranges + digit histograms
        "tile_table": ("hbm", 4.0 * P + 4.0 * (tiles + 200) + 8.0 * tiles),
        # 13 flop for every (pixel, list entry) pair the reference iterates (power + alpha test);
        # a contributing pair costs 22 flop in the forward and ~80 in the backward
        "blend_forward": ("fp32", 13.0 * (traversed - contributing) + 22.0 * contributing),
        "blend_backward": ("fp32", 13.0 * (traversed - contributing) + 80.0 * contributing),
        "memset_grad_acc": ("hbm", 48.0 * P),
        # fused loss: read both images, write the three derivative maps / read them + both images, write dL/dimage
        "ssim_forward": ("hbm", 20.0 * 3 * N), "ssim_backward": ("hbm", 24.0 * 3 * N),
        # p, g, m, v read + p, m, v written, 59 floats per Gaussian; one launch per STEP (reported per view)
        "adam_step": ("hbm", 28.0 * 59 * P / args.views_per_step),
        "preprocess_bwd": ("hbm", 563.0 * visible + (4.0 + 236.0) * (P - visible)),
    }
    return w, dict(num_rendered=R, visible=visible, traversed_pairs=traversed, contributing_pairs=contributing,
                   mean_list_len=R / tiles, pixels=N)


def kernel_table(report, views):
    rows = {}
    for line in report.strip().splitlines():
        name, n, ms = line.split()
        rows[name] = {"launches_per_view": int(n) / views, "ms_per_view": float(ms) / views}
    return rows


def roofline_entry(name, row, work, peaks, traffic=None):
    bound, amount = work[name]
    per_launch = amount / max(row["launches_per_view"], 1e-9)
    dur_s = row["ms_per_view"] / max(row["launches_per_view"], 1e-9) / 1e3
    if bound == "hbm":
        achieved, unit = per_launch / dur_s / 1e9, "GB/s"
        peak = peaks.get("hbm_gbs") or 6650.0
        src = "measured copy bandwidth (MEASURED_PEAKS.json)" if peaks.get("hbm_gbs") else "fallback 6.65 TB/s"
    else:
        achieved, unit = per_launch / dur_s / 1e12, "TFLOP/s"
        peak, src = FP32_PEAK_TFLOPS, "computed FP32 SIMT peak 148 SM x 128 FMA/clk x 1.965 GHz (MEASURED_PEAKS.json has no SIMT figure)"
    return {"kernel": name, "bound": bound, "achieved": round(achieved, 3), "peak": round(peak, 1), "unit": unit,
            "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": src,
            "avg_launch_ms": round(dur_s * 1e3, 4), "share_of_step": None}


# ----------------------------------------------------------------------------------------------- cpu baseline
def cpu_baseline(args):
    """C restatement (oracle/gsr_oracle.c, OpenMP) on the host cores: full forward + backward of ONE view of the same
    scene and camera set (a bounded sample of the 8-view step)."""
    import numpy as np
    import scenes
    from oracle import c_oracle
    c_oracle.build()
    cores = os.cpu_count() or 1
    c_oracle.set_threads(cores)
    sc = scenes.make_c3(P=args.gaussians)
    cam = scenes.c3_cameras(n=8, W=args.width, H=args.height)[0]
    S = c_oracle.Scene(W=cam.W, H=cam.H, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=np.zeros(3, np.float32),
                       means3D=sc.means3D.numpy(), opacities=sc.opacities.numpy(), viewmatrix=cam.viewmatrix.numpy(),
                       projmatrix=cam.projmatrix.numpy(), campos=cam.campos.numpy(), sh_degree=sc.sh_degree,
                       shs=sc.shs.numpy(), scales=sc.scales.numpy(), rotations=sc.rotations.numpy())
    dpix = np.random.default_rng(0).standard_normal((3, cam.H, cam.W)).astype(np.float32)
    n, t0 = 0, time.perf_counter()
    while True:
        fw = c_oracle.forward(S, want_invdepth=False)
        c_oracle.backward(S, fw, dpix, None)
        n += 1
        el = time.perf_counter() - t0
        if el > 10.0 or n >= 4:
            break
    return {"value": round(n / el, 4), "unit": "views/s", "cores": c_oracle.num_threads(), "kind": "port",
            "sample": f"{n} full fwd+bwd view(s) of the same {args.gaussians}-Gaussian scene at {args.width}x{args.height} "
                      f"(camera 0 of the step) by oracle/gsr_oracle.c with OpenMP, {el:.1f} s"}


# ----------------------------------------------------------------------------------------------- main
def main():
    # Libraries (NCCL prints its version banner) must not leak into the one-JSON-line contract: fd 1 goes to stderr
    # for the whole run and the result line is written to the saved descriptor at the end.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        return _main(real_stdout)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)


def _physical_gpu_index(local):
    """NVML enumerates all GPUs of the box; CUDA_VISIBLE_DEVICES may have re-numbered ours."""
    vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
    ids = [v.strip() for v in vis.split(",") if v.strip()]
    if local < len(ids) and ids[local].isdigit():
        return int(ids[local])
    return local


def emit(fd, obj):
    os.write(fd, (json.dumps(obj) + "\n").encode())


def _main(out_fd):
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference" and rank != 0:
        return 0
    import torch
    if not torch.cuda.is_available():
        if args.impl == "reference":
            cb = cpu_baseline(args)
            emit(out_fd, ({"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "views/s", "n_gpus": 0,
                              "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "cpu_baseline": cb,
                              "e2e": {"value": cb["value"], "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
            return 0
        raise SystemExit("bench.py needs a CUDA device: the rasterizer has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    import torch.distributed as dist
    bound_cpus = None
    if world > 1 and args.impl == "ours" and os.environ.get("GSR_NUMA_BIND", "1") != "0":
        from view_parallel import bind_rank_to_gpu_numa      # pinned host buffers next to the rank's GPU
        bound_cpus = bind_rank_to_gpu_numa(_physical_gpu_index(local))
    ref_world = 1 if args.impl == "reference" else world
    if ref_world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    import bench_extra
    if args.workload in ("c1", "c2", "slam", "knn"):
        # single-view configs: no view-parallel axis; under torchrun rank 0 alone runs them
        if ref_world > 1:
            dist.barrier()
        rc = 0
        if rank == 0:
            fn = bench_extra.run_knn if args.workload == "knn" else bench_extra.run_small
            rc = fn(args, dev, emit, out_fd, ClockSampler, local)
        if ref_world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return rc
    sc, cams, dpix_host, all_cams = build_workload(args, rank if ref_world > 1 else 0, ref_world, dev)
    if args.impl == "reference":
        from oracle import ref_cuda
        if not ref_cuda.available():
            cb = cpu_baseline(args)
            emit(out_fd, ({"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "views/s", "n_gpus": 1,
                              "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "cpu_baseline": cb,
                              "ms_per_step": None, "config": {"workload": "C3 (CPU port: reference CUDA build absent)"},
                              "e2e": {"value": cb["value"], "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
            return 0
        arm = ReferenceArm(args, sc, cams, dpix_host, dev)
    else:
        arm = OursArm(args, sc, cams, dpix_host, dev)
    exchange_kind = getattr(getattr(arm, "bucket", None), "exchange", "collective")
    fused_opt_used = bool(getattr(arm, "fused_opt", False))

    def barrier():
        if ref_world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    last_step_ms = []

    def timed(fn, steps, prof=False):
        """K steps bracketed by barrier + synchronize on both sides; device time by CUDA events, max over ranks."""
        barrier()
        if prof:
            arm._C.profile_enable(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        marks = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        t0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            fn()
            marks[i].record()
        e1.record()
        barrier()
        last_step_ms[:] = [a.elapsed_time(b) for a, b in zip([e0] + marks[:-1], marks)]
        wall = (time.perf_counter() - t0) * 1e3
        if prof:
            arm._C.profile_enable(False)
        ms = torch.tensor([e0.elapsed_time(e1), wall], device=dev, dtype=torch.float64)
        if ref_world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms[0]), float(ms[1])

    sampler = ClockSampler(local) if rank == 0 else None
    # warm-up: at least W steps, then until the step time has settled (the caching allocator needs a few different
    # num_rendered sizes before it stops calling cudaMalloc) -- at most 24 steps
    warm_run, prev = 0, None
    while warm_run < 24:
        t0 = time.perf_counter()
        arm.step()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        warm_run += 1
        settled = prev is not None and abs(dt - prev) <= 0.03 * prev
        prev = dt
        if ref_world > 1:                       # every rank must run the same number of steps (collective inside)
            flag = torch.tensor([1 if settled else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            settled = bool(int(flag[0]))
        if warm_run >= max(args.warmup, 3) and settled:
            break
    barrier()

    prof = args.impl == "ours" and not args.no_kernel_timing
    l0 = arm.launches()
    if rank == 0:
        sampler.start()
    ms_total, _wall = timed(arm.step, args.steps)
    launches = arm.launches() - l0
    step_ms = sorted(last_step_ms)
    # per-kernel device times: the same K steps again with one CUDA-event pair around every launch. Kept out of the
    # headline region because ~60 event records per view cost the host ~5 % there; kernel durations are unaffected.
    ms_prof = timed(arm.step, args.steps, prof=True)[0] if prof else None
    clocks = sampler.stop() if rank == 0 else None
    views_rank = len(cams)                                   # views this rank renders per step
    views_all = total_views(args, ref_world)                 # ... and all ranks together
    views = views_rank * args.steps
    ms_per_step = ms_total / args.steps
    value = views_all * args.steps / (ms_total / 1e3)
    parity = None
    if ref_world > 1 and args.impl == "ours" and args.workload == "raster" and not args.no_parity:
        arm.step()                                           # a fresh, un-profiled step: its exchanged bucket is what is checked
        torch.cuda.synchronize()
        parity = bench_extra.parity_multi_gpu(arm, args, all_cams, rank, ref_world, dev, rank_dpix)

    kernels, roof, roof_hbm, stats = None, None, None, None
    if prof:
        report = arm._C.profile_report()
        if report:
            kernels = kernel_table(report, views)
            work, stats = work_model(arm, args)
            peaks = {}
            try:
                peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            except Exception:
                pass
            ms_view = ms_per_step / max(views_rank, 1)
            traffic = {}
            try:
                traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
            except Exception:
                pass
            known = [k for k in kernels if k in work]
            top = max(known, key=lambda k: kernels[k]["ms_per_view"])
            roof = roofline_entry(top, kernels[top], work, peaks, traffic.get(top))
            roof["share_of_step"] = round(kernels[top]["ms_per_view"] / ms_view, 4)
            hbm = [k for k in known if work[k][0] == "hbm"]
            if hbm:
                th = max(hbm, key=lambda k: kernels[k]["ms_per_view"])
                roof_hbm = roofline_entry(th, kernels[th], work, peaks, traffic.get(th))
                roof_hbm["share_of_step"] = round(kernels[th]["ms_per_view"] / ms_view, 4)
            for k in kernels:
                kernels[k] = {a: round(b, 4) for a, b in kernels[k].items()}
                if k in work:
                    kernels[k]["frac_of_roofline"] = roofline_entry(k, kernels[k], work, peaks)["frac"] \
                        if kernels[k]["ms_per_view"] > 0 else None

    e2e = None
    if args.impl == "ours" and not args.no_e2e:
        arm.step_e2e()
        _dev_ms, wall_ms = timed(arm.step_e2e, args.steps)
        h2d, d2h = arm.e2e_bytes()
        e2e = {"value": round(views_all * args.steps / (wall_ms / 1e3), 3), "unit": "views/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "timing": "host wall clock around K steps incl. pinned H2D/D2H, max over ranks"}

    spec_stats = arm._C.speculation_stats() if args.impl == "ours" else None
    binding = arm._C.binding_name() if args.impl == "ours" else "reference pybind module (oracle/_ref/ref_dgr_C.so)"
    fp32_peak = None
    if rank == 0 and args.impl == "ours":
        try:
            fp32_peak = arm._C.fp32_peak_tflops(dev)
        except Exception:       # noqa: BLE001
            fp32_peak = None
    extras, cb = None, None
    if rank == 0 and world == 1 and args.impl == "ours":
        del arm
        torch.cuda.empty_cache()
        if args.workload == "raster" and not args.no_extras:
            extras = bench_extra.secondary(args, sc, cams, dpix_host, dev, OursArm, kernel_table, CHECK_LRS)
        if not args.no_cpu_baseline:
            cb = cpu_baseline(args)

    if rank == 0:
        kind = exchange_kind
        if fused_opt_used:
            kind = kind + "+adam"
        exchange_desc = {"multimem+adam": "gsr_exchange_adam_step launch per rank (gradient slice reduced through the NVSwitch multicast mapping, "
                                          "Adam on the slice, parameters multicast back)",
                         "peer+adam": "gsr_exchange_adam_step launch per rank (gradient slice from peer loads, Adam on the slice, "
                                      "parameters to every replica by peer stores)",
                         "multimem": "gsr_allreduce_sum launch per rank (NVSwitch multicast reduction over symmetric memory)",
                         "peer": "gsr_allreduce_sum launch per rank (NVLink peer loads/stores over symmetric memory)"}.get(
                             kind, "NCCL all-reduce")
        out = {"metric": METRIC, "value": round(value, 3), "unit": "views/s", "n_gpus": ref_world, "steps": args.steps,
               "warmup": args.warmup, "warmup_steps_run": warm_run, "ms_per_step": round(ms_per_step, 3), "higher_is_better": True,
               "scaling": args.scaling,
               "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "fwd_bwd_ms_per_iter": round(ms_per_step / max(views_rank, 1), 4),
               "config": {"workload": f"C3: {args.gaussians} Gaussians SH3 (Mip-NeRF360-scale mixture, seed 2), "
                                      f"{args.width}x{args.height}, orbit cameras; rasterizer forward+backward" +
                                      ("" if args.workload == "raster" else
                                       " + fused L1/SSIM loss per view + one Adam step over all parameters per step (full train-step)"),
                          "views_per_step_per_gpu": views_rank, "global_views_per_step": views_all,
                          "parallelism": f"view-parallel dp{ref_world}, replicated Gaussians" +
                                         (f", one {exchange_desc} of 59 floats/Gaussian per step" if ref_world > 1 else ""),
                          "l2": "inputs (708 MB parameters + 300 MB scratch per view) exceed the 126 MB L2; no explicit flush"},
               "step_ms_min_med_max": [round(step_ms[0], 3), round(step_ms[len(step_ms) // 2], 3), round(step_ms[-1], 3)],
               "clocks": clocks, "gpu_launches": launches, "binding": binding,
               "host": {"cpus_usable": len(os.sched_getaffinity(0)), "cpus_total": os.cpu_count(),
                        "rank0_bound_to_gpu_numa_cpus": (f"{bound_cpus[0]}-{bound_cpus[-1]} ({len(bound_cpus)})" if bound_cpus else None)},
               "kernel_timing_pass_ms_per_step": round(ms_prof / args.steps, 3) if ms_prof else None}
        if args.impl == "reference":
            out["impl"] = "reference"
            out["cpu_baseline"] = {"value": out["value"], "unit": "views/s", "cores": 0, "kind": "reference",
                                   "sample": "the reference has no CPU rasterizer: this arm is its own unmodified CUDA "
                                             "rasterizer (oracle/_ref/ref_dgr_C.so) on one B200, same step"}
            out["e2e"] = {"value": out["value"], "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
            out["gpu_launches"] = None
        else:
            out["e2e"] = e2e
            out["roofline"] = roof
            out["roofline_hbm"] = roof_hbm
            out["cpu_baseline"] = cb
            out["workload_stats"] = stats
            out["kernels"] = kernels
            if parity is not None:
                out["parity_multi_gpu"] = parity
            if extras:
                out.update(extras)
            if fp32_peak:
                out["fp32_simt_peak_measured_tflops"] = round(fp32_peak, 2)
                for r_ in (roof,):
                    if r_ and r_.get("bound") == "fp32":
                        r_["peak"] = round(fp32_peak, 2)
                        r_["frac"] = round(r_["achieved"] / fp32_peak, 4)
                        r_["peak_source"] = ("measured in this run: gsr_debug_fp32_peak, 16 independent FFMA chains per thread, "
                                             "8 x 256 threads per SM, CUDA events (MEASURED_PEAKS.json has no SIMT figure)")
                if kernels:
                    for k in ("blend_forward", "blend_backward"):
                        if k in kernels and kernels[k].get("frac_of_roofline") is not None:
                            kernels[k]["frac_of_roofline"] = round(kernels[k]["frac_of_roofline"] * FP32_PEAK_TFLOPS / fp32_peak, 4)
            try:
                out["speculation"] = dict(zip(("estimate_held", "repeated", "counted_first"), spec_stats))
            except Exception:       # noqa: BLE001
                pass
        emit(out_fd, out)
    if ref_world > 1:
        dist.destroy_process_group()
    if rank == 0 and parity is not None and not parity.get("ok", True):
        print("bench.py: parity_multi_gpu outside its bar (see the line's parity_multi_gpu field)", file=sys.stderr)
    if rank == 0 and parity is not None and parity.get("gross_mismatch", False):
        return 3                                             # the n-GPU gradients must equal the one-GPU ones
    return 0


if __name__ == "__main__":
    sys.exit(main())
