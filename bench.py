#!/usr/bin/env python
"""bench.py -- frames/s of the 5-camera Panoptic jln64 forward (heatmaps -> 3-D poses).

    python bench.py --gpus N --steps K --warmup W            # this package (CUDA hot path)
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU torch path (oracle port)

A step is one forward of the jln64 model from input heatmaps over one batch per GPU:
space-cube projection, CenterNet, NMS/top-k, z-column gather, C2CNet, proposal records, per-person
64^3 cube projection (10 per frame, bbox = 1.0 -> full cubes; records injected because random-init
weights give degenerate proposals), three max-planes, P2PNet, soft-argmax, WeightNet,
fusion.  The conv nets are PyTorch modules on both arms (north_star); everything else is this
package's CUDA kernels on ours and the torch-CPU restatement on the reference arm.

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from faster_voxelpose_b200 import synthetic as syn          # noqa: E402
from faster_voxelpose_b200.config import make_cfg            # noqa: E402

METRIC = "frames/s 5-cam Panoptic jln64"
UNIT = "frames/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="panoptic")
    ap.add_argument("--batch", type=int, default=32, help="frames per GPU per step")
    ap.add_argument("--heatmaps", default="G", choices=["G", "N"], help="Gaussian blobs or U[0,1) noise")
    ap.add_argument("--bbox", default="full", choices=["full", "crop"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="A/B: e2e driver launches eagerly instead of replaying CUDA graphs")
    ap.add_argument("--e2e-poses", action="store_true", help="also time the driver fed with 2-D poses (device-side splat)")
    ap.add_argument("--no-cudnn-benchmark", action="store_true", help="A/B: cuDNN heuristics instead of autotuning")
    ap.add_argument("--cached-grid", action="store_true", help="A/B: read reference-style cached sample grids")
    ap.add_argument("--no-tap-table", action="store_true", help="A/B: project every voxel in the kernel instead of reading the cached tap tables")
    ap.add_argument("--ref-frames", type=int, default=4, help="frames per step of the CPU reference arm")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons while the timed region runs: through NVML (the library behind
    nvidia-smi; ~1 ms per sample, so a 0.25 s region yields dozens) and, where NVML cannot be loaded, through
    the nvidia-smi binary itself (one sample per ~0.3 s)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self._stop_evt = index, [], threading.Event()
        self.source, self._nvml, self._handle = "nvidia-smi", None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            # LOCAL_RANK indexes CUDA_VISIBLE_DEVICES, NVML indexes the physical devices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self._nvml, self.source, self.index = pynvml, "nvml", phys
        except Exception:
            self._nvml = None

    def _sample_nvml(self):
        n, h = self._nvml, self._handle
        sm = n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)
        mx = n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM)
        r = n.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        bits = [n.nvmlClocksThrottleReasonHwSlowdown, n.nvmlClocksThrottleReasonHwThermalSlowdown,
                n.nvmlClocksThrottleReasonSwThermalSlowdown, n.nvmlClocksThrottleReasonSwPowerCap]
        return [str(sm), str(mx)] + ["Active" if r & b else "Not Active" for b in bits]

    def run(self):
        while not self._stop_evt.is_set():
            try:
                if self._nvml is not None:
                    self.samples.append(self._sample_nvml())
                else:
                    out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                    parts = [p.strip() for p in out.strip().split(",")]
                    if len(parts) >= 6:
                        self.samples.append(parts)
            except Exception:
                pass
            self._stop_evt.wait(0.01 if self._nvml is not None else 0.2)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        reasons = [n for i, n in enumerate(self.NAMES) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples), "source": self.source}


def build_inputs(cfg, cams, B, seed, family, bbox):
    V, J = len(cams), cfg.DATASET.NUM_JOINTS
    W, H = cfg.DATASET.HEATMAP_SIZE
    hm = syn.gaussian_heatmaps(seed, cfg, cams, B) if family == "G" else syn.noise_heatmaps(seed, B, V, J, H, W)
    props = syn.synthetic_proposals(seed + 1, cfg, B, bbox=bbox)
    return hm, props


def load_cams(workload):
    """Shipped calibration for the three dataset rigs; BASELINE config 4 (dense) runs on the Panoptic rig,
    config 5 (stress) on the seeded 31-camera dome."""
    if workload == "stress":
        return syn.dome_rig(31, seed=0)
    return syn.load_rig(workload if workload in ("panoptic", "shelf", "campus") else "panoptic")


def workload_name(args, cfg, V):
    P = cfg.CAPTURE_SPEC.MAX_PEOPLE
    n = cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS[0]
    return (f"{args.workload} jln64: batch {args.batch} frames/GPU, {V} cams, {cfg.DATASET.NUM_JOINTS}-joint "
            f"{cfg.DATASET.HEATMAP_SIZE[1]}x{cfg.DATASET.HEATMAP_SIZE[0]} heatmaps, "
            f"{'x'.join(str(v) for v in cfg.CAPTURE_SPEC.VOXELS_PER_AXIS)} space grid, {P} proposals x {n}^3 cubes")


# ------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the torch-CPU restatement of the reference (oracle/ref_torch.py)
# ------------------------------------------------------------------------------------------------
def run_cpu_reference(args, steps, warmup, frames):
    from faster_voxelpose_b200.lib.models import faster_voxelpose as m
    from oracle import ref_torch as RT

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = make_cfg(args.workload, device="cpu", CAPTURE_SPEC__MIN_SCORE=-1.0)
    cams = load_cams(args.workload)
    torch.manual_seed(0)
    model = m.get(cfg).eval()
    pipe = RT.RefTorchPipeline(cfg, model.pose_net.center_net, model.pose_net.c2c_net, model.joint_net.conv_net,
                               model.joint_net.weight_net)
    hm, props = build_inputs(cfg, cams, frames, 100, args.heatmaps, args.bbox)
    hm_t, props_t = torch.from_numpy(hm), torch.from_numpy(props)
    rt = torch.as_tensor(syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE), dtype=torch.float32)
    meta, cameras = {"seq": ["s"] * frames}, {"s": cams}
    for _ in range(warmup):
        pipe.forward(hm_t, meta, cameras, rt, proposals=props_t)
    t0 = time.perf_counter()
    for _ in range(steps):
        pipe.forward(hm_t, meta, cameras, rt, proposals=props_t)
    dt = time.perf_counter() - t0
    return {"value": frames * steps / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{frames} frame(s)/step x {steps} steps after {warmup} warm-up (sample grids cached, as the "
                      f"reference does); torch-CPU restatement oracle/ref_torch.py, torch {torch.__version__}",
            "ms_per_step": 1e3 * dt / steps}, cfg, len(cams)


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    base, cfg, V = run_cpu_reference(args, max(1, args.steps), max(1, args.warmup), args.ref_frames)
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": base["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args, cfg, V), "heatmaps": args.heatmaps, "bbox": args.bbox,
                   "sample_frames_per_step": args.ref_frames, "device": "host CPU"},
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def ours(args):
    import torch.distributed as dist

    from faster_voxelpose_b200 import _lib as L
    from faster_voxelpose_b200 import ops
    from faster_voxelpose_b200.lib.models import faster_voxelpose as m

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl=ours) needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # the reference's yaml (configs/panoptic/jln64.yaml:1-4, applied at run/validate.py:63-65)
    torch.backends.cudnn.benchmark = not args.no_cudnn_benchmark
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    cfg = make_cfg(args.workload, device=str(dev), CAPTURE_SPEC__MIN_SCORE=-1.0)
    cams = load_cams(args.workload)
    V, J, B = len(cams), cfg.DATASET.NUM_JOINTS, args.batch
    torch.manual_seed(0)
    model = m.get(cfg).eval().to(dev)
    for layer in (model.pose_net.project_layer, model.joint_net.project_layer):
        layer.use_cached_grid = bool(args.cached_grid)
        layer.use_tap_table = not args.no_tap_table
    hm_np, props_np = build_inputs(cfg, cams, B, 100 + rank, args.heatmaps, args.bbox)
    hm_host = torch.from_numpy(hm_np).pin_memory()
    hm = hm_host.to(dev)
    props = torch.from_numpy(props_np).to(dev)
    rt = torch.as_tensor(syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE), dtype=torch.float32,
                         device=dev)
    meta, cameras = {"seq": ["s"] * B}, {"s": cams}

    def step(x):
        with torch.no_grad():
            return model(meta=meta, input_heatmaps=x, cameras=cameras, resize_transform=rt, proposals=props)[0]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, n):
        barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(n):
            fn()
        e.record()
        barrier()
        ms = torch.tensor([s.elapsed_time(e)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(3, args.warmup)):
        step(hm)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    ops.TIMER.reset()
    ops.TIMER.enabled = True
    launches0 = ops.KERNEL_LAUNCHES[0]
    torch.cuda.profiler.start()  # no-op unless a profiler runs with --profile-from-start off
    ms = timed(lambda: step(hm), args.steps)
    torch.cuda.profiler.stop()
    launches = ops.KERNEL_LAUNCHES[0] - launches0
    ops.TIMER.enabled = False
    kern = ops.TIMER.summary()
    clocks = sampler.finish() if sampler else None

    # e2e: the public batched driver (runtime.StreamedInference).  Every step copies its 295 MB of pinned
    # host heatmaps to the device and its poses back inside the timed region; the copy of step i+1
    # overlaps the forward of step i (two staging buffers).
    e2e = e2e_poses = None
    if not args.no_e2e:
        from faster_voxelpose_b200.runtime import StreamedInference
        drv = StreamedInference(model, cameras, rt, dev, graph=not args.no_graph)
        sink = [0.0]

        def host_steps(n):
            for out in drv.run((hm_host for _ in range(n)), proposals=props):
                sink[0] += float(out[0, 0, 0, 0])          # touch the host result
        host_steps(3)
        n_e2e = max(3, args.steps)
        ms_e2e = timed(lambda: host_steps(n_e2e), 1)
        e2e = {"value": world * B * n_e2e / (ms_e2e * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(hm_host.numel() * 4), "d2h_bytes_per_step": int(B * cfg.CAPTURE_SPEC.MAX_PEOPLE * J * 5 * 4),
               "steps": n_e2e, "ms_per_step": ms_e2e / n_e2e,
               "api": "faster_voxelpose_b200.runtime.StreamedInference.run (H2D of step i+1 overlaps the forward of step i"
                      + ("" if args.no_graph else "; forward replayed as a CUDA graph per staging slot") + ")"}

        # Same driver fed with the 2-D poses the maps are made from (the reference's 'pred'/'gt' heatmap
        # sources): only the poses cross PCIe, ops.splat_heatmaps renders the maps in HBM (extra key, not the
        # contract's e2e: the step additionally contains the splat kernel).
        if args.e2e_poses:
            npeople = 10
            g = torch.Generator().manual_seed(1234 + rank)
            iw, ih = cfg.DATASET.IMAGE_SIZE
            root = torch.rand((B, V, npeople, 1, 2), generator=g) * torch.tensor([float(iw), float(ih)])
            joints = (root + 60.0 * torch.randn((B, V, npeople, J, 2), generator=g)).float().pin_memory()
            batch = {"joints": joints, "vis": None, "count": None}
            drv_p = StreamedInference(model, cameras, rt, dev, splat=dict(
                sigma=cfg.NETWORK.SIGMA, image_size=cfg.DATASET.IMAGE_SIZE, heatmap_size=cfg.DATASET.HEATMAP_SIZE))

            def pose_steps(n):
                for out in drv_p.run((batch for _ in range(n)), proposals=props):
                    sink[0] += float(out[0, 0, 0, 0])
            pose_steps(3)
            ms_p = timed(lambda: pose_steps(n_e2e), 1)
            e2e_poses = {"value": world * B * n_e2e / (ms_p * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(joints.numel() * 4),
                         "d2h_bytes_per_step": e2e["d2h_bytes_per_step"], "steps": n_e2e, "ms_per_step": ms_p / n_e2e,
                         "people_per_view": npeople,
                         "api": "StreamedInference.run on 2-D pose batches (fvp_splat_heatmaps renders the input maps on the device)"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # roofline of the dominant kernel (fused per-person projection + max-planes): algorithmic bytes per
    # launch = heatmaps of the frames it samples read once + planes written once (fused
    # variant).  The kernel is bound by the SM's 128 B/clk L1/shared data path (4 taps x 4*Jp bytes per
    # voxel-view), not by HBM: `onchip` reports that bound next to the HBM fraction the contract asks for.
    peak, peak_src = peaks()
    H, W = cfg.DATASET.HEATMAP_SIZE[1], cfg.DATASET.HEATMAP_SIZE[0]
    Jp = ops.padded_joints(J)
    roof = None
    k = kern.get("project_individual_planes")
    if k:
        tot_bytes, tap_bytes, taps, table_bytes = 0.0, 0.0, 0.0, 0.0
        uses_table = not args.no_tap_table and not args.cached_grid
        for info in k["info"]:
            frames = min(info["frames"], -(-info["cubes"] // cfg.CAPTURE_SPEC.MAX_PEOPLE))
            tot_bytes += 4.0 * V * J * H * W * frames + 4.0 * info["cubes"] * info["planes"]
            # cached tap records streamed once per voxel and view (the reference streams its 8-byte cached grid the same way)
            table_bytes += (20.0 if uses_table else 8.0 if args.cached_grid else 0.0) * V * info["cubes"] * info["voxels"]
            tap_bytes += 4.0 * 4 * Jp * V * info["cubes"] * info["voxels"]
            taps += 4.0 * V * J * info["cubes"] * info["voxels"]
        sec = k["ms_total"] * 1e-3
        achieved = tot_bytes / sec / 1e9
        sms = L.sm_count()
        l1_peak = sms * 128 * ((clocks or {}).get("sm_mhz") or 1965.0) * 1e6 / 1e9
        # dram__bytes_read+write of this launch from the committed ncu --set full capture (same shapes only)
        traffic = None
        prof = os.path.join(ROOT, "profiles", "r01_ncu_summary.json")
        if os.path.exists(prof) and args.workload == "panoptic" and B == 32 and args.heatmaps == "G" and args.bbox == "full":
            with open(prof) as fh:
                traffic = float(json.load(fh)["traffic_bytes_per_launch"])
        roof = {"kernel": "project_fused_kernel<Q,planes> (per-person cubes -> 3 max-planes, a12+a13)", "bound": "hbm",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": peak_src, "launches": k["launches"], "ms_avg": k["ms_avg"],
                "algorithmic_bytes_per_launch": tot_bytes / k["launches"], "gtaps_per_s": taps / sec / 1e9,
                "cached_positions": {"what": "tap-table (or cached-grid) bytes the launch streams on top of the algorithmic bytes: "
                                             "one record per voxel and view, no reuse; they are why `traffic` exceeds the algorithmic bytes",
                                     "bytes_per_launch": table_bytes / k["launches"],
                                     "achieved_incl": (tot_bytes + table_bytes) / sec / 1e9,
                                     "frac_incl": (tot_bytes + table_bytes) / sec / 1e9 / peak},
                "onchip": {"bound": "L1/shared data path, 128 B/clk/SM", "tap_bytes_per_launch": tap_bytes / k["launches"],
                           "achieved": tap_bytes / sec / 1e9, "peak": l1_peak, "unit": "GB/s",
                           "frac": tap_bytes / sec / 1e9 / l1_peak}}
    breakdown = {name: {"ms_per_step": v["ms_total"] / args.steps, "launches_per_step": v["launches"] / args.steps}
                 for name, v in kern.items()}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        base, _, _ = run_cpu_reference(args, steps=16, warmup=2, frames=args.ref_frames)   # ~10-12 s of CPU work
        cpu = {kk: base[kk] for kk in ("value", "unit", "cores", "kind", "sample")}

    line = {
        "metric": METRIC, "value": world * B * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args, cfg, V), "heatmaps": args.heatmaps, "bbox": args.bbox,
                   "parallelism": f"frames sharded x{world}, no data-path collective",
                   "l2_policy": f"inputs larger than L2 ({hm_host.numel() * 4 / 1e6:.0f} MB heatmaps + "
                                f"{hm_host.numel() * 4 * 2 * Jp / J * (H // 2 + 1) / (H / 2) / 1e6:.0f} MB row-paired copy + "
                                f"{3 * B * cfg.CAPTURE_SPEC.MAX_PEOPLE * Jp * cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS[0] ** 2 * 4 / 1e6:.0f} MB planes + "
                                "conv activations per step vs 126 MB L2)",
                   "proposals": "injected (SURVEY.md F6); detection still runs",
                   "conv_nets": "PyTorch/cuDNN fp32 (TF32 as torch defaults), BN folded, channels-last, cudnn.benchmark as in the reference yaml", "sample_positions": "cached sample grid (reference style, 8 B per view and voxel)" if args.cached_grid else
                   ("projected in the kernel" if args.no_tap_table else "cached tap tables (20 B per view and voxel, built once per rig)"),
                   "launch_mode": "value: eager launches (every library kernel is CUDA-event timed for the roofline); e2e: "
                                  + ("eager launches" if args.no_graph else "the same forward replayed as one CUDA graph per staging slot, "
                                     "so e2e can exceed value by the launch gaps it removes")},
        "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        "kernels": breakdown,
    }
    if not args.no_e2e and e2e_poses is not None:
        line["e2e_from_poses"] = e2e_poses
    _emit(line)
    if world > 1:
        dist.destroy_process_group()


def _emit(line: dict) -> None:
    """The one JSON line goes to the process's real stdout (see _quiet_stdout)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def _quiet_stdout() -> None:
    """Native libraries print to fd 1 (NCCL's version banner, cuDNN notes): point fd 1 at stderr for the
    run and keep a duplicate of the real stdout for the single JSON line."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


if __name__ == "__main__":
    a = parse()
    _quiet_stdout()
    if a.impl == "reference":
        reference_arm(a)
    else:
        ours(a)
