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
    ap.add_argument("--compact-proposals", action="store_true",
                    help="A/B: JointLocalizationNet compacts the valid proposal slots first (one host sync per forward)")
    ap.add_argument("--no-graph", action="store_true", help="A/B: e2e driver launches eagerly instead of replaying CUDA graphs")
    ap.add_argument("--e2e-poses", action="store_true", help="also time the driver fed with 2-D poses (device-side splat)")
    ap.add_argument("--no-cudnn-benchmark", action="store_true", help="A/B: cuDNN heuristics instead of autotuning")
    ap.add_argument("--cached-grid", action="store_true", help="A/B: read reference-style cached sample grids")
    ap.add_argument("--no-tap-table", action="store_true", help="A/B: project every voxel in the kernel instead of reading the cached tap tables")
    ap.add_argument("--ref-frames", type=int, default=4, help="frames per step of the CPU reference arm")
    ap.add_argument("--ref-port", action="store_true", help="CPU arm: run the torch restatement even when oracle/_ref is present")
    ap.add_argument("--e2e-depth", type=int, default=3, help="staging slots of the streamed driver (uploads in flight + 1)")
    ap.add_argument("--e2e-lanes", type=int, default=2, help="compute streams of the streamed driver: consecutive batches run on them in rotation")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra measurements (cube variant, noise heatmaps, reference on CUDA, z-slab)")
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
    """The reference's own eval forward on the host cores.  With oracle/_ref present (oracle/make_ref.py: the
    reference's UNMODIFIED lib/models, lib/core/proposal.py, lib/utils files) that is what runs (kind "reference");
    otherwise the bit-identical torch restatement oracle/ref_torch.py (kind "port")."""
    from oracle import ref_modules as RM

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = make_cfg(args.workload, device="cpu", CAPTURE_SPEC__MIN_SCORE=-1.0)
    cams = load_cams(args.workload)
    torch.manual_seed(0)
    if RM.available() and not args.ref_port:
        pipe = RM.RefModulePipeline(cfg, device="cpu")
        kind, what = "reference", ("the reference's own HumanDetectionNet + JointLocalizationNet modules (oracle/_ref, unmodified), "
                                   "composed as FasterVoxelPoseNet.forward composes them, proposal records injected between the two")
    else:
        from faster_voxelpose_b200.lib.models import faster_voxelpose as m
        from oracle import ref_torch as RT
        model = m.get(cfg).eval()
        pipe = RT.RefTorchPipeline(cfg, model.pose_net.center_net, model.pose_net.c2c_net, model.joint_net.conv_net,
                                   model.joint_net.weight_net)
        kind, what = "port", "torch-CPU restatement oracle/ref_torch.py (bit-identical to the reference modules) + this repo's conv nets"
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
    return {"value": frames * steps / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{frames} frame(s)/step x {steps} steps after {warmup} warm-up (sample grids cached, as the "
                      f"reference does) of the same workload (the GPU arm runs {args.batch} frames/step); {what}; torch {torch.__version__}",
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
# extra measurements of our arm (all CUDA-event timed on the launching stream, after warm-up)
# ------------------------------------------------------------------------------------------------
def _event_ms(fn, reps, dev, flush=None):
    """Median CUDA-event time of ``fn`` over ``reps`` runs; L2 is flushed before every run when ``flush`` is given."""
    ms = []
    for _ in range(reps):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize(dev)
        ms.append(s.elapsed_time(e))
    return float(np.median(ms))


def projection_extras(args, cfg, cams, model, hm, props, rt, dev):
    """Per-person projection on its own: honest work count from the windows, the same launch on noise heatmaps
    (nothing to skip), the interface-preserving cube-writing variant and its HBM fraction."""
    from faster_voxelpose_b200 import ops
    pl = model.joint_net.project_layer
    V, J, B = len(cams), cfg.DATASET.NUM_JOINTS, hm.shape[0]
    W, H = cfg.DATASET.HEATMAP_SIZE
    n = int(pl.consts.voxels[0])
    jp = ops.padded_joints(J)
    cameras = {"s": cams}
    rigs, geom = pl.rigs(cameras, V, dev), pl.geom(V, rt)
    frame_rig = rigs.frame_rig(["s"] * B)
    p7 = props.reshape(-1, 7).float().contiguous()
    N = p7.shape[0]
    frame = torch.arange(B, dtype=torch.int32, device=dev).repeat_interleave(N // B)
    win, _ = ops.individual_windows(p7, None, frame, pl.consts)
    w = win.cpu().numpy().astype(np.int64)
    ok = w[:, 9] == 0
    live = int(np.prod(w[ok, 6:9] - w[ok, 3:6], axis=1).sum())
    hm_rp = ops.rowpair(hm, cache=False)
    nz_frac = float(hm_rp.occ[:, 0].sum().item()) / (B * V * 2 * (H // 2 + 1) * W)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    planes = torch.empty((3, N, n, n, jp), dtype=torch.float32, device=dev)
    out = (planes[0], planes[1], planes[2])
    run = lambda src: pl.project_planes(src, frame_rig, rigs, geom, p7, frame, out=out, nhwc=True)
    run(hm_rp)
    ms_here = _event_ms(lambda: run(hm_rp), 5, dev, flush)
    noise = torch.from_numpy(syn.noise_heatmaps(7, B, V, J, H, W)).to(dev)
    noise_rp = ops.rowpair(noise, cache=False)
    del noise
    run(noise_rp)
    ms_noise = _event_ms(lambda: run(noise_rp), 5, dev, flush)
    del noise_rp, planes
    res = {"cubes": N, "voxels_per_cube": n ** 3, "live_voxels": live, "voxels_in_window_frac": live / (N * n ** 3),
           "nonzero_pixel_records_frac": nz_frac, "ms_alone_l2_flushed": ms_here,
           "ms_noise_heatmaps": ms_noise, "noise_note": "same launch on U[0,1) heatmaps: no all-zero pixel record, nothing is skipped"}
    # interface-preserving variant: what ProjectLayer.forward (the drop-in) runs -- cubes written to HBM, then planes3
    cube_bytes = 4.0 * N * J * n ** 3
    free, _ = torch.cuda.mem_get_info(dev)
    if cube_bytes * 1.2 < free:
        cubes = torch.empty((N, J, n, n, n), dtype=torch.float32, device=dev)
        grid = pl.cached_grids(rigs, geom, dev)
        proj = lambda: ops.project_cubes(hm_rp, rigs.dev, frame_rig, geom, pl.axes(dev), [n, n, n], N, win=win,
                                         cached_grid=grid, mode=pl.mode, out=cubes)
        proj()
        ms_cube = _event_ms(proj, 3, dev, flush)
        ms_p3 = _event_ms(lambda: ops.planes3(cubes), 3, dev, flush)
        peak, _ = peaks()
        alg = 4.0 * V * J * H * W * B + cube_bytes
        res["cube_variant"] = {
            "what": "fvp_project_cubes for the same windows (the cube interface of ProjectLayer.forward): every voxel of "
                    "every cube is written, zeros outside the window", "ms": ms_cube, "bytes_written": cube_bytes,
            "algorithmic_bytes": alg, "achieved_gbs": alg / (ms_cube * 1e-3) / 1e9, "frac_of_hbm_peak": alg / (ms_cube * 1e-3) / 1e9 / peak,
            "planes3_ms": ms_p3, "planes3_gbs": (cube_bytes + 3.0 * 4 * N * J * n * n) / (ms_p3 * 1e-3) / 1e9}
        del cubes
    return res


def reference_cuda_extra(args, cfg, cams, model, hm, props, rt, dev, our_fps):
    """The reference's own modules (oracle/_ref) on torch-CUDA, same inputs, same weights, grids cached."""
    from oracle import ref_modules as RM
    if not RM.available():
        return {"unavailable": "oracle/_ref not present (oracle/make_ref.py needs /root/reference at build time)"}
    B = hm.shape[0]
    pipe = RM.RefModulePipeline(cfg, device=dev)
    pipe.load_weights_from(model)
    meta, cameras = {"seq": ["s"] * B}, {"s": cams}
    fwd = lambda: pipe.forward(hm, meta, cameras, rt, proposals=props)
    for _ in range(2):
        out = fwd()
    torch.cuda.synchronize(dev)
    ms = _event_ms(fwd, 3, dev)
    with torch.no_grad():
        ours = model(meta=meta, input_heatmaps=hm, cameras=cameras, resize_transform=rt, proposals=props)[0]
    dev_mm = float((ours[..., :3] - out[0][..., :3]).abs().max())
    del pipe
    torch.cuda.empty_cache()
    return {"value": B / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "frames_per_step": B,
            "what": "the reference's HumanDetectionNet + JointLocalizationNet (unmodified, oracle/_ref) on torch-CUDA "
                    "on this GPU, same heatmaps / proposals / weights, sample grids cached, CUDA-event timed",
            "ours_over_reference_cuda": our_fps / (B / (ms * 1e-3)), "max_joint_deviation_mm": dev_mm}


def zslab_extra(dev, rank, world):
    """BASELINE config 4 through the product path: HumanDetectionNet with the 256x256x64 space grid split into z
    slabs over all ranks (same 4 frames on every rank), plane combine by NCCL / peer pull / NVLS.  Every rank calls
    this; the times are the max over ranks."""
    import torch.distributed as dist
    from faster_voxelpose_b200 import ops
    from faster_voxelpose_b200.lib.models.human_detection_net import HumanDetectionNet
    B = 4
    cfg = make_cfg("dense", device=str(dev), CAPTURE_SPEC__MIN_SCORE=-1.0)
    cams = syn.load_rig("panoptic")
    torch.manual_seed(0)
    net = HumanDetectionNet(cfg).eval().to(dev)
    hm = torch.from_numpy(syn.gaussian_heatmaps(5, cfg, cams, B)).to(dev)          # the same frames on every rank
    rt = torch.as_tensor(syn.resize_transform(cfg.DATASET.ORI_IMAGE_SIZE, cfg.DATASET.IMAGE_SIZE), dtype=torch.float32, device=dev)
    meta, cameras = {"seq": ["s"] * B}, {"s": cams}
    pl = net.project_layer
    rigs, geom = pl.rigs(cameras, len(cams), dev), pl.geom(len(cams), rt)
    fr = rigs.frame_rig(meta["seq"])
    hm_rp = ops.rowpair(hm, cache=False)

    def timed_max(fn, reps):
        for _ in range(3):
            fn()
        dist.barrier()
        torch.cuda.synchronize(dev)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(reps):
            fn()
        e.record()
        torch.cuda.synchronize(dev)
        t = torch.tensor([s.elapsed_time(e) / reps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    with torch.no_grad():
        plane_ref = net.space_plane(hm_rp, fr, rigs, geom).clone()
        ms_plane_1 = timed_max(lambda: net.space_plane(hm_rp, fr, rigs, geom), 20)
        want = net(hm, meta, cameras, rt)
        ms_net_1 = timed_max(lambda: net(hm, meta, cameras, rt), 10)
        modes, best = {}, None
        for mode in ("nccl", "pull", "multimem"):
            try:
                net.enable_zshard(combine=mode)
                same = True
                for _ in range(3):
                    same = same and bool(torch.equal(net.space_plane(hm_rp, fr, rigs, geom), plane_ref))
                ms = timed_max(lambda: net.space_plane(hm_rp, fr, rigs, geom), 20)
            except Exception as exc:
                modes[mode] = {"unavailable": str(exc)[:160]}
                continue
            flag = torch.tensor([int(same)], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            modes[mode] = {"ms_plane": ms, "bit_identical": bool(flag.item())}
            if best is None or ms < modes[best]["ms_plane"]:
                best = mode
        net.enable_zshard(combine=best)
        got = net(hm, meta, cameras, rt)
        same_net = torch.tensor([int(all(torch.equal(a, b) for a, b in zip(got, want)))], device=dev)
        dist.all_reduce(same_net, op=dist.ReduceOp.MIN)
        ms_net_z = timed_max(lambda: net(hm, meta, cameras, rt), 10)
        # what the combine costs with nothing to combine: the device-side barriers of the symmetric-memory handle
        barrier_ms = None
        peers = [p for p in net._peer.values()]
        if peers:
            h = peers[0].handle
            barrier_ms = timed_max(lambda: h.barrier(channel=0), 50)
    net.disable_zshard()
    return {"workload": f"dense 256x256x64 space grid over 8000x8000x2000 mm (BASELINE config 4), {B} frames replicated on every rank, "
                        f"z slabs of {-(-64 // world)} over {world} ranks, through HumanDetectionNet.enable_zshard",
            "ms_unsharded_1gpu": ms_plane_1, "ms_sharded": modes[best]["ms_plane"], "combine": best, "modes": modes,
            "bit_identical": all(m.get("bit_identical", True) for m in modes.values()) and bool(same_net.item()),
            "detection_net_ms": {"unsharded": ms_net_1, "z_sharded": ms_net_z, "outputs_bit_identical": bool(same_net.item())},
            "barrier_only_ms": barrier_ms, "timing": "CUDA events, max over ranks, 20 launches after 3 warm-up"}


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
    affinity = None
    if world > 1 and hasattr(os, "sched_setaffinity"):
        # one disjoint block of host cores per rank: the launch thread, the clock sampler and the first touch of the
        # pinned staging buffers stay off the other ranks' cores (all GPUs of the box hang off one NUMA node)
        cpus = sorted(os.sched_getaffinity(0))
        per = max(1, len(cpus) // world)
        mine = cpus[local * per:(local + 1) * per] or cpus
        try:
            os.sched_setaffinity(0, mine)
            affinity = [mine[0], mine[-1]]
        except OSError:
            affinity = None
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

    # The timed forward runs JointLocalizationNet over all B x MAX_PEOPLE proposal slots with the invalid ones masked on the device
    # (its `static_shapes` mode: fixed shapes, no host synchronisation -- what the e2e driver's CUDA graphs replay); the other
    # mode compacts the valid slots first, which costs one host sync per forward (--compact-proposals for the A/B).  Every slot
    # of the bench's proposals is valid, so both modes do the same work.
    model.joint_net.static_shapes = not args.compact_proposals

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
        drv = StreamedInference(model, cameras, rt, dev, graph=not args.no_graph, depth=args.e2e_depth, lanes=args.e2e_lanes)
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
                      + ("" if args.no_graph else "; forward replayed as a CUDA graph per staging slot")
                      + (f"; consecutive steps on {drv.lanes} compute streams in rotation" if drv.lanes > 1 else "")
                      + ")"}

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

    # z-sharded dense grid through the product path (every rank takes part; BASELINE config 4)
    zslab = None
    if world > 1 and not args.no_extras:
        try:
            zslab = zslab_extra(dev, rank, world)
        except Exception as exc:  # reported, never fatal for the headline line
            zslab = {"error": f"{type(exc).__name__}: {str(exc)[:300]}"}

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # roofline of the dominant kernel (fused per-person projection + max-planes): algorithmic bytes per
    # launch = heatmaps of the frames it samples read once + planes written once (fused
    # variant).  The work is counted from the windows: only voxels inside [start, end) of a non-skipped
    # window are projected (the bench's proposals are uniform in a 2 m high space, so about 63 % of the
    # 320 x 64^3 voxels).  The kernel is bound by the SM's L1 data path, not by HBM: `onchip` reports that
    # bound next to the HBM fraction the contract asks for.
    peak, peak_src = peaks()
    H, W = cfg.DATASET.HEATMAP_SIZE[1], cfg.DATASET.HEATMAP_SIZE[0]
    Jp = ops.padded_joints(J)
    roof = None
    k = kern.get("project_individual_planes")
    extras = None
    if k and not args.no_extras:
        try:
            extras = projection_extras(args, cfg, cams, model, hm, props, rt, dev)
        except Exception as exc:
            extras = {"error": f"{type(exc).__name__}: {str(exc)[:300]}"}
    if k:
        uses_table = not args.no_tap_table and not args.cached_grid
        sec = k["ms_total"] * 1e-3
        nl = k["launches"]
        frames = sum(min(i["frames"], -(-i["cubes"] // cfg.CAPTURE_SPEC.MAX_PEOPLE)) for i in k["info"]) / nl
        planes_bytes = sum(4.0 * i["cubes"] * i["planes"] for i in k["info"]) / nl
        alg = 4.0 * V * J * H * W * frames + planes_bytes
        achieved = alg / (sec / nl) / 1e9
        all_vox = sum(i["cubes"] * i["voxels"] for i in k["info"]) / nl
        live = extras.get("live_voxels") if extras and "live_voxels" in extras else None
        vox = live if live is not None else all_vox
        sms = L.sm_count()
        clk = ((clocks or {}).get("sm_mhz") or 1965.0) * 1e6
        # L1 data path for a gather: every voxel-view needs two 128-byte pixel records; a request that touches
        # several lines is served at 2.07 cycles per line (measured: constant-weight build of the kernel, 2.09)
        floor_ms = vox * V * 2 * 2.07 / (sms * clk) * 1e3
        # dram__bytes of this launch from the committed ncu --set full capture; only quoted while the capture still
        # describes this kernel (same shapes, duration within 10 % of the live launch)
        traffic, traffic_note = None, "no ncu capture committed for this kernel build"
        prof = os.path.join(ROOT, "profiles", "r02_ncu_summary.json")
        if os.path.exists(prof) and args.workload == "panoptic" and B == 32 and args.heatmaps == "G" and args.bbox == "full":
            with open(prof) as fh:
                cap = json.load(fh)
            if abs(cap["duration_ms"] - k["ms_avg"]) <= 0.10 * k["ms_avg"]:
                traffic, traffic_note = float(cap["traffic_bytes_per_launch"]), f"from {cap['source']}"
            else:
                traffic_note = (f"capture in profiles/r02_ncu_summary.json is stale ({cap['duration_ms']:.2f} ms there, "
                                f"{k['ms_avg']:.2f} ms now): re-run tools/profile_round.sh")
        roof = {"kernel": "project_walk3_kernel (per-person cubes -> 3 max-planes, a12+a13; fvp_project_walk.cu)", "bound": "hbm",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_note": traffic_note, "peak_source": peak_src, "launches": nl, "ms_avg": k["ms_avg"],
                "algorithmic_bytes_per_launch": alg,
                "work": {"cubes": all_vox / (cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS[0] ** 3), "voxels_in_window": vox,
                         "voxels_in_window_frac": vox / all_vox, "counted_from_windows": live is not None,
                         "voxel_views_in_window": vox * V,
                         "nonzero_pixel_records_frac": (extras or {}).get("nonzero_pixel_records_frac")},
                "gtaps_per_s": 4.0 * V * J * vox / (sec / nl) / 1e9,
                "cached_positions": {"what": "tap-table bytes the launch streams on top of the algorithmic bytes: one 4-byte record per "
                                             "voxel-view in the window, plus 16 bytes of weights for the samples that are gathered",
                                     "bytes_per_launch_upper": (20.0 if uses_table else 8.0 if args.cached_grid else 0.0) * V * vox},
                "onchip": {"bound": "L1 data path of a gather: 2 x 128-byte pixel records per voxel-view, 2.07 cycles per line in a "
                                    "multi-line request (measured; DESIGN.md section 4)",
                           "floor_ms_if_every_sample_is_gathered": floor_ms, "achieved_ms": k["ms_avg"],
                           "frac": floor_ms / k["ms_avg"],
                           "note": "above 1 when samples over all-zero pixel records are skipped (occupancy bitmap); "
                                   "ms_noise_heatmaps in `projection` is the same launch with nothing to skip"}}
        if extras:
            roof["projection"] = extras
    breakdown = {name: {"ms_per_step": v["ms_total"] / args.steps, "launches_per_step": v["launches"] / args.steps}
                 for name, v in kern.items()}
    # the HBM-bound kernel beside the projection: the tcgen05 tail of P2PNet reads three 32-channel tensors of the person planes
    # and writes the J output maps
    if roof is not None and "res_tail_tf32" in kern and kern["res_tail_tf32"]["launches"]:
        n_ind = cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS[0]
        planes = 3 * B * cfg.CAPTURE_SPEC.MAX_PEOPLE
        alg_tail = float(planes * n_ind * n_ind * (3 * 32 + J) * 4)
        ms_tail = kern["res_tail_tf32"]["ms_total"] / kern["res_tail_tf32"]["launches"]
        roof["other_kernels"] = {"res_tail_tf32": {
            "what": f"res_tail_kernel: {planes} planes of {n_ind} x {n_ind}, 3 x 32 channels in, {J} maps out",
            "bound": "hbm", "ms_avg": ms_tail, "algorithmic_bytes_per_launch": alg_tail,
            "achieved_gbs": alg_tail / (ms_tail * 1e-3) / 1e9, "frac": alg_tail / (ms_tail * 1e-3) / 1e9 / peak}}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        base, _, _ = run_cpu_reference(args, steps=16, warmup=2, frames=args.ref_frames)   # ~10-12 s of CPU work
        cpu = {kk: base[kk] for kk in ("value", "unit", "cores", "kind", "sample")}

    ref_cuda = None
    if world == 1 and not args.no_extras:
        try:
            ref_cuda = reference_cuda_extra(args, cfg, cams, model, hm, props, rt, dev, B * args.steps / (ms * 1e-3))
        except Exception as exc:
            ref_cuda = {"error": f"{type(exc).__name__}: {str(exc)[:300]}"}

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
                   "proposal_slots": "valid slots compacted first (one host sync per forward)" if args.compact_proposals
                   else "all B x MAX_PEOPLE slots, invalid ones masked on the device (no host sync; all slots of this workload are valid)",
                   "launch_mode": "value: eager launches (every library kernel is CUDA-event timed for the roofline); e2e: "
                                  + ("eager launches" if args.no_graph else "the same forward replayed as one CUDA graph per staging slot, "
                                     "so e2e can exceed value by the launch gaps it removes")},
        "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        "kernels": breakdown, "reference_cuda": ref_cuda,
    }
    if zslab is not None:
        line["zslab"] = zslab
    if affinity is not None:
        line["config"]["host_affinity"] = f"rank-local cores {affinity[0]}-{affinity[1]} (disjoint block per rank)"
    line["config"]["e2e_staging_slots"] = args.e2e_depth
    line["config"]["e2e_compute_lanes"] = args.e2e_lanes
    if not args.no_e2e and e2e_poses is not None:
        line["e2e_from_poses"] = e2e_poses
    _emit(line)
    if world > 1:
        dist.barrier()
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
