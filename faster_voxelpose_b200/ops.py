"""Tensor-level wrappers over the C ABI (one function per entry point) plus the small host-side
state the modules share: packed camera rigs, projection constants and the channels-last copy
of the heatmaps.

Nothing here computes on the CPU: every function launches CUDA kernels from libfvp_b200.so on the
current torch stream and raises if the tensors are not on a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import weakref
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib as L

__all__ = [
    "pack_cameras", "RigTable", "make_geom", "axis_tables", "channels_last", "rowpair", "rowpair_for_planes", "is_split17", "sample_grid",
    "individual_windows", "project_cubes", "project_planes", "topk_windows", "planes3", "nms_topk", "gather_topk", "proposal_records",
    "softargmax", "fuse", "weightnet_features", "padded_joints", "KERNEL_LAUNCHES", "TIMER",
]

# running count of kernels launched through this module (bench.py reports it as gpu_launches)
KERNEL_LAUNCHES = [0]


def _launched(n: int = 1) -> None:
    KERNEL_LAUNCHES[0] += n


class KernelTimer:
    """Optional CUDA-event timing of individual launches (bench.py's roofline numbers).  Events are
    recorded on the launching stream; ``summary()`` synchronises and returns per-kernel stats."""

    def __init__(self):
        self.enabled = False
        self.records: Dict[str, list] = {}

    def start(self, name: str, device, **info):
        if not self.enabled:
            return None
        s = torch.cuda.Event(enable_timing=True)
        e = torch.cuda.Event(enable_timing=True)
        s.record(torch.cuda.current_stream(device))
        self.records.setdefault(name, []).append((s, e, info))
        return e, device

    @staticmethod
    def stop(tok):
        if tok is not None:
            tok[0].record(torch.cuda.current_stream(tok[1]))

    def reset(self):
        self.records = {}

    def summary(self):
        torch.cuda.synchronize()
        out = {}
        for name, recs in self.records.items():
            ms = [s.elapsed_time(e) for s, e, _ in recs]
            out[name] = {"launches": len(ms), "ms_total": float(sum(ms)), "ms_avg": float(sum(ms) / len(ms)),
                         "info": [r[2] for r in recs]}
        return out


TIMER = KernelTimer()


def padded_joints(J: int) -> int:
    return 4 * ((int(J) + 3) // 4)


# ----------------------------------------------------------------------------- cameras
def pack_cameras(cams: Sequence[Dict]) -> np.ndarray:
    """[V,24] fp32 records R9 T3 fx fy cx cy k3 p2 pad3; float64 -> fp32 exactly like the
    reference's unfold_camera_param (lib/utils/cameras.py:11-18)."""
    out = np.zeros((len(cams), L.CAM_STRIDE), dtype=np.float32)
    for v, c in enumerate(cams):
        out[v, 0:9] = np.asarray(c["R"], dtype=np.float64).reshape(9).astype(np.float32)
        out[v, 9:12] = np.asarray(c["T"], dtype=np.float64).reshape(3).astype(np.float32)
        out[v, 12:16] = np.array([c["fx"], c["fy"], c["cx"], c["cy"]], dtype=np.float64).astype(np.float32)
        out[v, 16:19] = np.asarray(c["k"], dtype=np.float64).reshape(3).astype(np.float32)
        out[v, 19:21] = np.asarray(c["p"], dtype=np.float64).reshape(2).astype(np.float32)
    return out


def _cam_list(entry) -> List[Dict]:
    """cameras[seq] is a list, or an int-keyed dict (reference lib/dataset/shelf.py:147-152)."""
    if isinstance(entry, dict):
        return [entry[k] for k in sorted(entry.keys())]
    return list(entry)


class RigTable:
    """Device table [R][V][24] of every sequence in a ``cameras`` dict plus the per-batch
    ``frame_rig`` index vector.  Replaces the reference's per-sequence ``sample_grid`` cache key
    (lib/models/project_whole.py:144-159)."""

    def __init__(self, cameras: Dict, num_views: int, device: torch.device, packed=None):
        self.seqs = list(cameras.keys())
        self.index = {s: i for i, s in enumerate(self.seqs)}
        if packed is None:
            packed = []
            for s in self.seqs:
                cams = _cam_list(cameras[s])
                assert len(cams) == num_views, "inconsistent number of cameras"
                packed.append(pack_cameras(cams))
        self.num_views = num_views
        self.host = np.stack(packed) if packed else np.zeros((0, num_views, L.CAM_STRIDE), np.float32)
        self.dev = torch.from_numpy(self.host).to(device)
        self._frame_rig: Dict[Tuple, torch.Tensor] = {}

    def frame_rig(self, seqs: Sequence) -> torch.Tensor:
        key = tuple(seqs)
        t = self._frame_rig.get(key)
        if t is None:
            for s in key:
                assert s in self.index, "missing camera parameters for the current sequence"
            t = torch.tensor([self.index[s] for s in key], dtype=torch.int32, device=self.dev.device)
            if len(self._frame_rig) > 256:
                self._frame_rig.clear()
            self._frame_rig[key] = t
        return t


# ----------------------------------------------------------------------------- constants
def make_geom(cfg, num_views: int, resize_transform) -> L.Geom:
    """Pack ProjectLayer's constants (project_whole.py:29-32,100-113) for the kernels."""
    ds = cfg.DATASET
    w, h = int(ds.HEATMAP_SIZE[0]), int(ds.HEATMAP_SIZE[1])
    g = L.Geom()
    g.V, g.J, g.H, g.W = int(num_views), int(ds.NUM_JOINTS), h, w
    g.clamp_max = float(max(ds.ORI_IMAGE_SIZE[0], ds.ORI_IMAGE_SIZE[1]))
    if isinstance(resize_transform, torch.Tensor):
        t = resize_transform.detach().to("cpu", torch.float32).numpy().reshape(6)
    else:
        t = np.asarray(resize_transform, dtype=np.float64).astype(np.float32).reshape(6)
    for i in range(6):
        g.t[i] = float(t[i])
    g.hm_w, g.hm_h = float(w), float(h)
    g.img_w, g.img_h = float(ds.IMAGE_SIZE[0]), float(ds.IMAGE_SIZE[1])
    g.wm1, g.hm1 = float(w - 1), float(h - 1)
    return g


def axis_tables(space_size, space_center, voxels) -> List[torch.Tensor]:
    """Per-axis voxel-centre tables, computed with torch.linspace on the host exactly like the
    reference's compute_grid (project_whole.py:62-70) so the values carry torch's own rounding."""
    return [torch.linspace(-space_size[a] / 2, space_size[a] / 2, int(voxels[a])) + space_center[a]
            for a in range(3)]


# ----------------------------------------------------------------------------- heatmap layout
_cl_cache: Dict[str, object] = {"ref": None, "version": -1, "out": None}


def channels_last(heatmaps: torch.Tensor, cache: bool = True) -> torch.Tensor:
    """[B,V,J,H,W] -> channels-last [B,V,H,W,Jp].  The last conversion is remembered per source
    tensor object (and its version counter), so the per-frame calls the reference interface makes
    (joint_localization_net.py:148-154) convert once per batch."""
    hm = L.require_cuda(heatmaps, "heatmaps")
    B, V, J, H, W = hm.shape
    if cache and torch.cuda.is_current_stream_capturing():
        cache = False
    if cache:
        ref = _cl_cache["ref"]
        if ref is not None and ref() is heatmaps and _cl_cache["version"] == heatmaps._version:
            return _cl_cache["out"]
    Jp = padded_joints(J)
    out = torch.empty((B, V, H, W, Jp), dtype=torch.float32, device=hm.device)
    L.check(L.load().fvp_heatmaps_channels_last(L.ptr(hm), B * V, J, H, W, Jp, L.ptr(out),
                                                L.stream_handle(hm.device)), "fvp_heatmaps_channels_last")
    _launched()
    if cache:
        _cl_cache.update(ref=weakref.ref(heatmaps), version=heatmaps._version, out=out)
    return out


_rp_cache: Dict[str, object] = {"ref": None, "version": -1, "out": None}
_rp17_cache: Dict[str, object] = {"ref": None, "version": -1, "out": None}


def rowpair(heatmaps: torch.Tensor, cache: bool = True, occupancy: bool = True, split17: bool = False) -> torch.Tensor:
    """[B,V,J,H,W] -> row-paired [B,V,2,H/2+1,W,2*Jp]: the layout the projection kernels gather from
    (north|south rows adjacent per channel quad, one 256-bit load each; include/fvp_b200.h).  With ``occupancy``
    the pre-pass also records which pixel records are all zero (``.occ`` attribute of the result, uint32
    [B, fvp_occupancy_words]); ``project_planes`` hands it to the kernel, which then skips samples whose taps are
    all +0 -- same bits out, far fewer loads for heatmaps rendered from 2-D poses.  The last conversion is
    remembered per source tensor object and version, so the per-frame calls of the reference interface
    (joint_localization_net.py:148-154) convert once per batch; never while a CUDA graph is being captured (the
    conversion must be part of the graph).

    ``split17`` (17-joint heatmaps only): the split 16 + 1 layout of fvp_heatmaps_rowpair_split17 -- an opaque [B, n]
    buffer (36 floats per record position: the 16-channel records, then one float4 per position for channel 16) tagged
    ``.split17 = (B, 20)``.  Only ``project_planes`` with a tap table reads it (the row-walking kernel); it is what
    the model's forward uses for Shelf / Campus, see ``rowpair_for_planes``."""
    hm = L.require_cuda(heatmaps, "heatmaps")
    B, V, J, H, W = hm.shape
    if split17 and J != 17:
        raise L.FvpError(f"rowpair(split17=True) is the layout of 17-joint heatmaps (got {J} joints)")
    if cache and torch.cuda.is_current_stream_capturing():
        cache = False
    store = _rp17_cache if split17 else _rp_cache
    if cache:
        ref = store["ref"]
        if ref is not None and ref() is heatmaps and store["version"] == heatmaps._version:
            return store["out"]
    Jp = padded_joints(J)
    occ = None
    if occupancy:
        words = int(L.load().fvp_occupancy_words(V, H, W))
        occ = torch.empty((B, words), dtype=torch.int32, device=hm.device)
    if split17:
        out = torch.empty((B, V * 2 * (H // 2 + 1) * W * 36), dtype=torch.float32, device=hm.device)   # shape[0] = frames; not frame-major inside
        L.check(L.load().fvp_heatmaps_rowpair_split17(L.ptr(hm), B, V, H, W, L.ptr(out), L.ptr(occ),
                                                      L.stream_handle(hm.device)), "fvp_heatmaps_rowpair_split17")
        out.split17 = (B, Jp)
        _launched(2 if occupancy else 1)
    else:
        out = torch.empty((B, V, 2, H // 2 + 1, W, 2 * Jp), dtype=torch.float32, device=hm.device)
        if occupancy:
            L.check(L.load().fvp_heatmaps_rowpair_occ(L.ptr(hm), B, V, J, H, W, Jp, L.ptr(out), L.ptr(occ),
                                                      L.stream_handle(hm.device)), "fvp_heatmaps_rowpair_occ")
            _launched(2)   # memset + kernel
        else:
            L.check(L.load().fvp_heatmaps_rowpair(L.ptr(hm), B * V, J, H, W, Jp, L.ptr(out), L.stream_handle(hm.device)),
                    "fvp_heatmaps_rowpair")
            _launched()
    if occupancy:
        out.occ = occ
    if cache:
        store.update(ref=weakref.ref(heatmaps), version=heatmaps._version, out=out)
    return out


def rowpair_for_planes(heatmaps: torch.Tensor, cache: bool = True, layers=(), num_rigs: int = 1) -> torch.Tensor:
    """The row-paired layout for callers that only take fused planes (the model's forward): the split 16 + 1 layout
    for 17-joint heatmaps when every projection layer in ``layers`` will read tap tables (so that the row-walking
    kernel serves them), else ``rowpair``."""
    split = (heatmaps.shape[2] == 17 and L.get_option(L.OPT_WALK_VARIANT) >= 0 and len(layers) > 0
             and all(layer.tap_tables_fit(num_rigs, int(heatmaps.shape[1])) for layer in layers))
    return rowpair(heatmaps, cache=cache, split17=split)


def is_split17(hm_rp: torch.Tensor) -> bool:
    return getattr(hm_rp, "split17", None) is not None


def _staged_dims(hm_rp: torch.Tensor):
    """(B, Jp) of a row-paired heatmap tensor."""
    tag = getattr(hm_rp, "split17", None)
    if tag is not None:
        return int(tag[0]), int(tag[1])
    if hm_rp.dim() != 6 or hm_rp.shape[2] != 2:
        raise L.FvpError("expected row-paired heatmaps from ops.rowpair(): [B,V,2,H/2+1,W,2*Jp]")
    return int(hm_rp.shape[0]), int(hm_rp.shape[5]) // 2


# ----------------------------------------------------------------------------- kernels
def sample_grid(cams_dev: torch.Tensor, geom: L.Geom, axes: Sequence[torch.Tensor]) -> torch.Tensor:
    """The reference's cached grid for one rig: [V, X*Y*Z, 2]."""
    X, Y, Z = [int(a.numel()) for a in axes]
    out = torch.empty((geom.V, X * Y * Z, 2), dtype=torch.float32, device=cams_dev.device)
    L.check(L.load().fvp_sample_grid(L.ptr(cams_dev), C.byref(geom), L.ptr(axes[0]), L.ptr(axes[1]), L.ptr(axes[2]),
                                     X, Y, Z, L.ptr(out), L.stream_handle(cams_dev.device)), "fvp_sample_grid")
    _launched()
    return out


def tap_table(cams_dev: torch.Tensor, geom: L.Geom, axes: Sequence[torch.Tensor]) -> torch.Tensor:
    """Tap table of one rig: [V, 20 * npad] bytes (npad float4 weights, then npad packed records, per view)."""
    X, Y, Z = [int(a.numel()) for a in axes]
    npad = (X * Y * Z + 3) // 4 * 4
    out = torch.empty((geom.V, npad * L.TAP_ENTRY_BYTES), dtype=torch.uint8, device=cams_dev.device)
    L.check(L.load().fvp_tap_table(L.ptr(cams_dev), C.byref(geom), L.ptr(axes[0]), L.ptr(axes[1]), L.ptr(axes[2]),
                                   X, Y, Z, L.ptr(out), L.stream_handle(cams_dev.device)), "fvp_tap_table")
    _launched()
    return out


def _grid_flags(cached_grid: Optional[torch.Tensor]) -> int:
    """A float tensor is the reference-style sample grid, a byte tensor the tap table of tap_table()."""
    if cached_grid is None:
        return 0
    return L.FLAG_TAP_TABLE if cached_grid.dtype == torch.uint8 else L.FLAG_CACHED_GRID


def individual_windows(props: torch.Tensor, valid: Optional[torch.Tensor], frame: torch.Tensor, consts
                       ) -> Tuple[torch.Tensor, torch.Tensor]:
    """props [N,7] -> (win [N,12] int32, offset [N,3]); ``consts`` is an IndividualConsts."""
    props = L.require_cuda(props, "proposal_centers")
    N = props.shape[0]
    win = torch.empty((N, L.WIN_STRIDE), dtype=torch.int32, device=props.device)
    offset = torch.empty((N, 3), dtype=torch.float32, device=props.device)
    if N:
        if valid is not None:
            valid = L.require_cuda(valid, "valid", torch.uint8)
        frame = L.require_cuda(frame, "frame", torch.int32)
        L.check(L.load().fvp_individual_windows(
            L.ptr(props), L.ptr(valid), L.ptr(frame), N, L.f3(consts.scale), L.f3(consts.bias), L.i3(consts.fine),
            L.f3(consts.whole_size), L.f3(consts.ind_size), L.i3(consts.voxels), L.ptr(win), L.ptr(offset),
            L.stream_handle(props.device)), "fvp_individual_windows")
        _launched()
    return win, offset


def project_cubes(hm_cl: torch.Tensor, rig_dev: torch.Tensor, frame_rig: torch.Tensor, geom: L.Geom,
                  axes: Sequence[torch.Tensor], dims: Sequence[int], num_cubes: int,
                  win: Optional[torch.Tensor] = None, cached_grid: Optional[torch.Tensor] = None,
                  mode: int = L.MODE_DIRECT, out: Optional[torch.Tensor] = None, use_occupancy: bool = True) -> torch.Tensor:
    """Fused projection/gather/mean/clamp: returns cubes [C,J,nx,ny,nz] (``hm_cl``: ops.rowpair(); its occupancy
    words, when present, let the kernel skip samples over all-zero pixel records -- same cubes)."""
    B, Jp = _staged_dims(hm_cl)
    if is_split17(hm_cl):
        raise L.FvpError("project_cubes reads the packed row-paired layout; rowpair(split17=True) is for project_planes")
    nx, ny, nz = [int(d) for d in dims]
    if out is None:
        out = torch.empty((num_cubes, geom.J, nx, ny, nz), dtype=torch.float32, device=hm_cl.device)
    flags = mode | _grid_flags(cached_grid)
    if num_cubes:
        tok = TIMER.start("project_individual" if win is not None else "project_whole", hm_cl.device,
                          cubes=int(num_cubes), voxels=nx * ny * nz, frames=int(B))
        occ = getattr(hm_cl, "occ", None) if use_occupancy else None
        L.check(L.load().fvp_project_cubes_occ(
            L.ptr(hm_cl), B, Jp, L.ptr(rig_dev), L.ptr(frame_rig), C.byref(geom), L.ptr(axes[0]), L.ptr(axes[1]),
            L.ptr(axes[2]), int(axes[0].numel()), int(axes[1].numel()), int(axes[2].numel()), L.ptr(cached_grid),
            num_cubes, L.ptr(win), nx, ny, nz, L.ptr(out), flags, L.ptr(occ), L.stream_handle(hm_cl.device)),
            "fvp_project_cubes")
        TIMER.stop(tok)
        _launched()
    return out


def project_planes(hm_cl: torch.Tensor, rig_dev: torch.Tensor, frame_rig: torch.Tensor, geom: L.Geom,
                   axes: Sequence[torch.Tensor], dims: Sequence[int], num_cubes: int,
                   win: Optional[torch.Tensor] = None, cached_grid: Optional[torch.Tensor] = None,
                   mode: int = L.MODE_DIRECT, want=(True, True, True), zslab: Optional[Tuple[int, int]] = None,
                   out=None, cubes: Optional[torch.Tensor] = None, nhwc: bool = False, use_occupancy: bool = True):
    """Fused projection + orthographic max-planes (the cube never goes to HBM unless ``cubes`` is
    given): returns (xy [C,J,nx,ny], xz [C,J,nx,zs], yz [C,J,ny,zs]) with zs = z1 - z0 of ``zslab``
    (default the whole cube); entries not wanted are None.  With ``nhwc`` the planes are stored
    channels-last with the joints padded to Jp ([C,n1,n2,Jp] in memory) and returned as [C,Jp,n1,n2]
    views in torch's channels_last memory format -- what the conv nets consume without a copy.
    ``hm_cl`` is the row-paired heatmap tensor of ops.rowpair(); its occupancy words (``hm_cl.occ``), when present,
    let the kernel skip samples over all-zero pixel records (``use_occupancy=False``: A/B switch, same results)."""
    B, Jp = _staged_dims(hm_cl)
    nx, ny, nz = [int(d) for d in dims]
    z0, z1 = (0, nz) if zslab is None else (int(zslab[0]), int(zslab[1]))
    dev = hm_cl.device
    J = geom.J
    zs = z1 - z0
    shapes = [(num_cubes, nx, ny, Jp), (num_cubes, nx, zs, Jp), (num_cubes, ny, zs, Jp)] if nhwc else \
             [(num_cubes, J, nx, ny), (num_cubes, J, nx, zs), (num_cubes, J, ny, zs)]
    if out is not None:
        xy, xz, yz = out
        for t, shp in zip((xy, xz, yz), shapes):
            if t is not None and (tuple(t.shape) != shp or not t.is_contiguous() or t.dtype != torch.float32):
                raise L.FvpError(f"project_planes: out tensor must be contiguous fp32 of shape {shp}")
    else:
        xy, xz, yz = [torch.empty(shp, dtype=torch.float32, device=dev) if w else None for shp, w in zip(shapes, want)]
    flags = mode | _grid_flags(cached_grid) | (L.FLAG_PLANES_NHWC if nhwc else 0) | (L.FLAG_SPLIT17 if is_split17(hm_cl) else 0)
    if num_cubes:
        tok = TIMER.start("project_individual_planes" if win is not None else "project_whole_planes", dev,
                          cubes=int(num_cubes), voxels=nx * ny * (z1 - z0), frames=int(B),
                          planes=sum(int(t.numel()) for t in (xy, xz, yz) if t is not None) // int(num_cubes))
        occ = getattr(hm_cl, "occ", None) if use_occupancy else None
        L.check(L.load().fvp_project_planes_occ(
            L.ptr(hm_cl), B, Jp, L.ptr(rig_dev), L.ptr(frame_rig), C.byref(geom), L.ptr(axes[0]), L.ptr(axes[1]),
            L.ptr(axes[2]), int(axes[0].numel()), int(axes[1].numel()), int(axes[2].numel()), L.ptr(cached_grid),
            num_cubes, L.ptr(win), nx, ny, nz, z0, z1, L.ptr(cubes), L.ptr(xy), L.ptr(xz), L.ptr(yz), flags,
            L.ptr(occ), L.stream_handle(dev)), "fvp_project_planes")
        TIMER.stop(tok)
        _launched(1 + sum(t is not None for t in (xy, xz, yz)))   # the kernel + one memset per plane
    elif nhwc:
        for t in (xy, xz, yz):
            if t is not None:
                t.zero_()
    if nhwc and out is None:
        xy, xz, yz = [t.permute(0, 3, 1, 2) if t is not None else None for t in (xy, xz, yz)]
    return xy, xz, yz


def planes_combine_peers(xy_ptrs: Sequence[int], xz_ptrs: Optional[Sequence[int]], yz_ptrs: Optional[Sequence[int]],
                         zstart: Sequence[int], n_xy: int, rows_xz: int, rows_yz: int, out_xy: torch.Tensor,
                         out_xz: Optional[torch.Tensor], out_yz: Optional[torch.Tensor]) -> None:
    """One kernel: out_xy = max over the ranks' partial xy planes, out_xz / out_yz = the ranks' slabs joined
    along z, every input read through a (peer) device pointer.  ``*_ptrs`` are raw device addresses, one per
    rank (symmetric-memory ``buffer_ptrs`` plus an offset); ``zstart`` has R + 1 entries."""
    R = len(xy_ptrs)
    arr = lambda ptrs: None if ptrs is None else (C.c_void_p * R)(*[int(p) for p in ptrs])
    zs = (C.c_int32 * (R + 1))(*[int(z) for z in zstart])
    dev = out_xy.device
    for t, name in ((out_xy, "out_xy"), (out_xz, "out_xz"), (out_yz, "out_yz")):
        if t is not None:
            L.require_cuda(t, name)
            if not t.is_contiguous():
                raise L.FvpError(f"planes_combine_peers: {name} must be contiguous")
    L.check(L.load().fvp_planes_combine_peers(arr(xy_ptrs), arr(xz_ptrs), arr(yz_ptrs), R, zs, int(n_xy), int(rows_xz),
                                             int(rows_yz), L.ptr(out_xy), L.ptr(out_xz), L.ptr(out_yz),
                                             L.stream_handle(dev)), "fvp_planes_combine_peers")
    _launched()


def planes_combine_multimem(mc_xy_partial: int, mc_xy_full: int, xz_slab: Optional[torch.Tensor], mc_xz_full: int,
                            yz_slab: Optional[torch.Tensor], mc_yz_full: int, world: int, rank: int,
                            zstart: Sequence[int], n_xy: int, rows_xz: int, rows_yz: int, device) -> None:
    """NVLS form of planes_combine_peers: ``mc_*`` are multicast addresses (symmetric-memory ``multicast_ptr`` plus
    an offset); this rank reduces and broadcasts its share of the xy plane and broadcasts its xz / yz slabs."""
    zs = (C.c_int32 * (world + 1))(*[int(z) for z in zstart])
    L.check(L.load().fvp_planes_combine_multimem(
        int(mc_xy_partial), int(mc_xy_full), L.ptr(xz_slab), int(mc_xz_full) if xz_slab is not None else None,
        L.ptr(yz_slab), int(mc_yz_full) if yz_slab is not None else None, int(world), int(rank), zs, int(n_xy),
        int(rows_xz), int(rows_yz), L.stream_handle(torch.device(device))), "fvp_planes_combine_multimem")
    _launched()


def topk_windows(flat: torch.Tensor, X: int, Y: int, Z: int) -> torch.Tensor:
    """flat [B,P] top-k cells -> [B*P,12] int32 column windows (one 1x1xZ cube per proposal)."""
    flat = L.require_cuda(flat, "flat", torch.int64)
    B, P = flat.shape
    win = torch.empty((B * P, L.WIN_STRIDE), dtype=torch.int32, device=flat.device)
    L.check(L.load().fvp_topk_windows(L.ptr(flat), B, P, int(X), int(Y), int(Z), L.ptr(win),
                                      L.stream_handle(flat.device)), "fvp_topk_windows")
    _launched()
    return win


def planes3(cubes: torch.Tensor, want=(True, True, True), out=None):
    """Orthographic max-planes of cubes [C,J,nx,ny,nz] -> (xy, xz, yz) (None where not wanted).
    ``out``: optional triple of preallocated contiguous tensors (e.g. slices of one [3,N,J,n,n]
    buffer that feeds P2PNet directly)."""
    cubes = L.require_cuda(cubes, "cubes")
    Cn, J, nx, ny, nz = cubes.shape
    dev = cubes.device
    if out is not None:
        xy, xz, yz = out
        for t, shp in ((xy, (Cn, J, nx, ny)), (xz, (Cn, J, nx, nz)), (yz, (Cn, J, ny, nz))):
            if t is not None and (tuple(t.shape) != shp or not t.is_contiguous() or t.dtype != torch.float32):
                raise L.FvpError(f"planes3: out tensor must be contiguous fp32 of shape {shp}")
    else:
        xy = torch.empty((Cn, J, nx, ny), dtype=torch.float32, device=dev) if want[0] else None
        xz = torch.empty((Cn, J, nx, nz), dtype=torch.float32, device=dev) if want[1] else None
        yz = torch.empty((Cn, J, ny, nz), dtype=torch.float32, device=dev) if want[2] else None
    if Cn:
        tok = TIMER.start("planes3", dev, cubes=int(Cn), voxels=nx * ny * nz)
        L.check(L.load().fvp_planes3(L.ptr(cubes), Cn * J, nx, ny, nz, L.ptr(xy), L.ptr(xz), L.ptr(yz),
                                     L.stream_handle(dev)), "fvp_planes3")
        TIMER.stop(tok)
        _launched()
    return xy, xz, yz


def nms_topk(prob_map: torch.Tensor, max_num: int):
    """[B,1,X,Y] -> (values [B,P], index2d [B,P,2] int64, flat [B,P] int64)."""
    pm = L.require_cuda(prob_map, "prob_map")
    B, X, Y = pm.shape[0], pm.shape[-2], pm.shape[-1]
    dev = pm.device
    scratch = torch.empty((B, X * Y), dtype=torch.float32, device=dev)
    vals = torch.empty((B, max_num), dtype=torch.float32, device=dev)
    idx2 = torch.empty((B, max_num, 2), dtype=torch.int64, device=dev)
    flat = torch.empty((B, max_num), dtype=torch.int64, device=dev)
    L.check(L.load().fvp_nms_topk(L.ptr(pm), B, X, Y, int(max_num), L.ptr(scratch), L.ptr(vals), L.ptr(idx2), L.ptr(flat),
                                  L.stream_handle(dev)), "fvp_nms_topk")
    _launched()
    return vals, idx2, flat


def gather_topk(bbox: Optional[torch.Tensor], cubes: Optional[torch.Tensor], flat: torch.Tensor):
    """bbox [B,2,X,Y] -> [B,P,2]; cubes [B,J,X,Y,Z] -> [B,P,J,Z] at the top-k cells."""
    flat = L.require_cuda(flat, "flat", torch.int64)
    B, P = flat.shape
    dev = flat.device
    mb = cols = None
    J, Z = 1, 1
    if bbox is not None:
        bbox = L.require_cuda(bbox, "bbox")
        X, Y = bbox.shape[-2:]
        mb = torch.empty((B, P, 2), dtype=torch.float32, device=dev)
    if cubes is not None:
        cubes = L.require_cuda(cubes, "cubes")
        J, X, Y, Z = cubes.shape[1:]
        cols = torch.empty((B, P, J, Z), dtype=torch.float32, device=dev)
    L.check(L.load().fvp_gather_topk(L.ptr(bbox), L.ptr(cubes), L.ptr(flat), B, P, J, X, Y, Z, L.ptr(mb), L.ptr(cols),
                                     L.stream_handle(dev)), "fvp_gather_topk")
    _launched()
    return mb, cols


def proposal_records(idx2: torch.Tensor, conf2d: torch.Tensor, hm1d: torch.Tensor, match_bbox: torch.Tensor,
                     scale, bias, min_score: float) -> torch.Tensor:
    """-> proposal_centers [B,P,7] (eval-mode ProposalLayer)."""
    idx2 = L.require_cuda(idx2, "topk_index", torch.int64)
    conf2d = L.require_cuda(conf2d, "topk_confs")
    hm1d = L.require_cuda(hm1d, "proposal_heatmaps_1d")
    match_bbox = L.require_cuda(match_bbox, "match_bbox_preds")
    B, P = conf2d.shape
    rec = torch.empty((B, P, 7), dtype=torch.float32, device=conf2d.device)
    L.check(L.load().fvp_proposal_records(L.ptr(idx2), L.ptr(conf2d), L.ptr(hm1d), L.ptr(match_bbox), B * P,
                                          int(hm1d.shape[-1]), L.f3(scale), L.f3(bias), float(min_score), L.ptr(rec),
                                          L.stream_handle(rec.device)), "fvp_proposal_records")
    _launched()
    return rec


def softargmax(feat: torch.Tensor, center_grid: torch.Tensor, beta: float):
    """feat [3,N,J,n,n] -> (pose [3,N,J,2], maxprob [3,N,J])."""
    feat = L.require_cuda(feat, "joint_features")
    grid = L.require_cuda(center_grid, "center_grid")
    _, N, J = feat.shape[:3]
    M = int(feat.shape[3] * feat.shape[4]) if feat.dim() == 5 else int(feat.shape[3])
    pose = torch.empty((3, N, J, 2), dtype=torch.float32, device=feat.device)
    maxprob = torch.empty((3, N, J), dtype=torch.float32, device=feat.device)
    if N:
        L.check(L.load().fvp_softargmax(L.ptr(feat), L.ptr(grid), N, J, M, float(beta), L.ptr(pose), L.ptr(maxprob),
                                        L.stream_handle(feat.device)), "fvp_softargmax")
        _launched()
    return pose, maxprob


def fuse(pose: torch.Tensor, maxprob: torch.Tensor, offset: torch.Tensor, weights: torch.Tensor):
    """Adds offsets to ``pose`` in place; returns (fused [N,J,3], conf [N])."""
    _, N, J = pose.shape[:3]
    weights = L.require_cuda(weights, "weights").reshape(3, N, J)
    offset = L.require_cuda(offset, "offset")
    fused = torch.empty((N, J, 3), dtype=torch.float32, device=pose.device)
    conf = torch.empty((N,), dtype=torch.float32, device=pose.device)
    if N:
        L.check(L.load().fvp_fuse(L.ptr(pose), L.ptr(maxprob), L.ptr(offset), L.ptr(weights), N, J, L.ptr(fused),
                                  L.ptr(conf), L.stream_handle(pose.device)), "fvp_fuse")
        _launched()
    return fused, conf


def weightnet_features(maps: torch.Tensor, w_host: np.ndarray, b_host: np.ndarray, tf32: bool = False) -> torch.Tensor:
    """maps [M,n,n] -> [M,F]: WeightNet's conv3x3 + folded BN + max-pool + ReLU + global mean in one
    kernel.  ``w_host`` [F,9] / ``b_host`` [F] are contiguous fp32 numpy arrays (kernel parameters).
    ``tf32``: run the convolution on TF32 tensor cores (n % 8 == 0), as torch's own Conv2d does by default."""
    maps = L.require_cuda(maps, "joint_features")
    M, n = int(maps.shape[0]), int(maps.shape[-1])
    F = int(b_host.shape[0])
    out = torch.empty((M, F), dtype=torch.float32, device=maps.device)
    if M:
        tok = TIMER.start("weightnet_features", maps.device, maps=M)
        fn = L.load().fvp_weightnet_features_tf32 if tf32 else L.load().fvp_weightnet_features
        L.check(fn(L.ptr(maps), M, n, w_host.ctypes.data, b_host.ctypes.data, F,
                   L.ptr(out), L.stream_handle(maps.device)), "fvp_weightnet_features")
        TIMER.stop(tok)
        _launched()
    return out


def _nhwc(x: torch.Tensor, what: str) -> torch.Tensor:
    """A [N,C,H,W] fp32 CUDA tensor whose memory is dense [N][H][W][C] (made so if it is not)."""
    x = L.require_cuda(x, what, contiguous=False)
    if x.dim() != 4 or x.shape[1] % 4:
        raise L.FvpError(f"{what}: expected [N,C,H,W] with C % 4 == 0, got {tuple(x.shape)}")
    if not x.permute(0, 2, 3, 1).is_contiguous():
        x = x.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)
    return x


def maxpool2_nhwc(x: torch.Tensor) -> torch.Tensor:
    """MaxPool2d(2) on a channels-last activation [N,C,H,W] -> channels-last [N,C,H//2,W//2]."""
    x = _nhwc(x, "activation")
    N, C, H, W = (int(v) for v in x.shape)
    out = torch.empty((N, H // 2, W // 2, C), dtype=torch.float32, device=x.device).permute(0, 3, 1, 2)
    if out.numel():
        tok = TIMER.start("conv_glue", x.device)
        L.check(L.load().fvp_maxpool2_nhwc(L.ptr(x), N, H, W, C, L.ptr(out), L.stream_handle(x.device)),
                "fvp_maxpool2_nhwc")
        TIMER.stop(tok)
        _launched()
    return out


def upshuffle_add(y: torch.Tensor, s: torch.Tensor) -> torch.Tensor:
    """y [N,4C,H,W] (1x1 form of a k=2/s=2 transposed conv, channels ordered (a,b,c)) pixel-shuffled to
    [N,C,2H,2W] plus the skip tensor ``s``; everything channels-last."""
    y, s = _nhwc(y, "upsampled"), _nhwc(s, "skip")
    N, C4, H, W = (int(v) for v in y.shape)
    C = C4 // 4
    if tuple(s.shape) != (N, C, 2 * H, 2 * W):
        raise L.FvpError(f"upshuffle_add: skip {tuple(s.shape)} does not match {(N, C, 2 * H, 2 * W)}")
    out = torch.empty((N, 2 * H, 2 * W, C), dtype=torch.float32, device=y.device).permute(0, 3, 1, 2)
    if out.numel():
        tok = TIMER.start("conv_glue", y.device)
        L.check(L.load().fvp_upshuffle_add(L.ptr(y), L.ptr(s), N, H, W, C, L.ptr(out), L.stream_handle(y.device)),
                "fvp_upshuffle_add")
        TIMER.stop(tok)
        _launched()
    return out


def pointwise_planar(a: torch.Tensor, up: Optional[torch.Tensor], w_host: np.ndarray, b_host: np.ndarray,
                     resid: Optional[torch.Tensor] = None, resid_bias: Optional[np.ndarray] = None) -> torch.Tensor:
    """1x1 conv of the channels-last activation ``a`` [N,32,H,W] (+ the un-shuffled up-sampling ``up``
    [N,128,H/2,W/2]) to contiguous planar maps [N,Cout,H,W].  ``w_host`` [Cout,32], ``b_host`` [Cout].
    With ``resid`` / ``resid_bias`` the activation is formed on the way in as ``relu(a + resid_bias + resid)``: ``a`` is
    then the second convolution of a residual block without its epilogue, ``resid`` the block's skip input."""
    a = _nhwc(a, "activation")
    N, C, H, W = (int(v) for v in a.shape)
    up_ptr = None
    if up is not None:
        up = _nhwc(up, "upsampled")
        if tuple(up.shape) != (N, 4 * C, H // 2, W // 2) or H % 2 or W % 2:
            raise L.FvpError(f"pointwise_planar: up {tuple(up.shape)} does not match {(N, 4 * C, H // 2, W // 2)}")
        up_ptr = L.ptr(up)
    if (resid is None) != (resid_bias is None):
        raise L.FvpError("pointwise_planar: resid and resid_bias go together")
    cout = int(b_host.shape[0])
    out = torch.empty((N, cout, H, W), dtype=torch.float32, device=a.device)
    if out.numel():
        tok = TIMER.start("conv_glue", a.device)
        if resid is None:
            L.check(L.load().fvp_pointwise_planar(L.ptr(a), up_ptr, N, H, W, C, w_host.ctypes.data, b_host.ctypes.data,
                                                  cout, L.ptr(out), L.stream_handle(a.device)), "fvp_pointwise_planar")
        else:
            resid = _nhwc(resid, "residual")
            rb = np.ascontiguousarray(resid_bias, dtype=np.float32)
            if tuple(resid.shape) != (N, C, H, W) or rb.shape != (C,):
                raise L.FvpError(f"pointwise_planar: resid {tuple(resid.shape)} / bias {rb.shape} do not match {(N, C, H, W)}")
            L.check(L.load().fvp_pointwise_planar_res(L.ptr(a), L.ptr(resid), rb.ctypes.data, up_ptr, N, H, W, C,
                                                      w_host.ctypes.data, b_host.ctypes.data, cout, L.ptr(out),
                                                      L.stream_handle(a.device)), "fvp_pointwise_planar_res")
        TIMER.stop(tok)
        _launched()
    return out


def splat_heatmaps(joints: torch.Tensor, vis: Optional[torch.Tensor], count: Optional[torch.Tensor], sigma: float,
                   image_size: Sequence[int], heatmap_size: Sequence[int]) -> torch.Tensor:
    """Input heatmaps from 2-D poses (reference ``JointsDataset.generate_input_heatmap``, augmentation off).

    joints [B,V,N,J,2] fp32 network-input pixels; vis [B,V,N,J] uint8 or None; count [B,V] int32 (people per
    view) or None.  ``image_size`` / ``heatmap_size`` are (w, h) as in the reference's cfg.
    Returns [B,V,J,H,W] fp32."""
    joints = L.require_cuda(joints, "joints")
    if joints.dim() != 5 or joints.shape[-1] != 2:
        raise L.FvpError(f"joints must be [B,V,N,J,2], got {tuple(joints.shape)}")
    B, V, N, J, _ = (int(v) for v in joints.shape)
    if vis is not None:
        vis = L.require_cuda(vis, "joints_vis", dtype=torch.uint8)
        if tuple(vis.shape) != (B, V, N, J):
            raise L.FvpError(f"joints_vis must be {(B, V, N, J)}, got {tuple(vis.shape)}")
    if count is not None:
        count = L.require_cuda(count, "people count", dtype=torch.int32)
        if tuple(count.shape) != (B, V):
            raise L.FvpError(f"people count must be {(B, V)}, got {tuple(count.shape)}")
    iw, ih = (int(v) for v in image_size)
    hw, hh = (int(v) for v in heatmap_size)
    out = torch.empty((B, V, J, hh, hw), dtype=torch.float32, device=joints.device)
    if out.numel():
        tok = TIMER.start("splat_heatmaps", joints.device)
        L.check(L.load().fvp_splat_heatmaps(L.ptr(joints), L.ptr(vis) if vis is not None else None,
                                            L.ptr(count) if count is not None else None, B * V, N, J, float(sigma),
                                            iw, ih, hw, hh, L.ptr(out), L.stream_handle(joints.device)),
                "fvp_splat_heatmaps")
        TIMER.stop(tok)
        _launched()
    return out


def c2c_forward(columns: torch.Tensor, weights: torch.Tensor, op_table: np.ndarray) -> torch.Tensor:
    """C2CNet (eval) on z-columns [n,J,20] -> [n,1,20] in one kernel.  ``weights``: flat fp32 device buffer,
    ``op_table``: contiguous int32 host array [nops,10] (see ``fvp_c2c_forward`` in include/fvp_b200.h)."""
    columns = L.require_cuda(columns, "z-columns")
    weights = L.require_cuda(weights, "c2c weights")
    n, J, Z = (int(v) for v in columns.shape)
    out = torch.empty((n, 1, Z), dtype=torch.float32, device=columns.device)
    if n:
        tok = TIMER.start("c2c_forward", columns.device, columns=n)
        L.check(L.load().fvp_c2c_forward(L.ptr(columns), n, J, Z, L.ptr(weights), op_table.ctypes.data,
                                         int(op_table.shape[0]), L.ptr(out), L.stream_handle(columns.device)),
                "fvp_c2c_forward")
        TIMER.stop(tok)
        _launched()
    return out


def conv7_fragments(weight: torch.Tensor) -> torch.Tensor:
    """[16,16,7,7] or [16,20,7,7] filters -> the B-fragment buffer of ``fvp_conv7x7_c16_tf32`` / ``_c20_`` (TF32-rounded,
    ties away): channels 0..15 as two k8 steps per tap, channels 16..19 (20-channel inputs) as one k4 step per tap."""
    if tuple(weight.shape) not in ((16, 16, 7, 7), (16, 20, 7, 7)):
        raise L.FvpError(f"conv7_fragments: expected [16,16,7,7] or [16,20,7,7] filters, got {tuple(weight.shape)}")
    w = weight.detach().float().contiguous()
    w = ((w.view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)
    wt = w[:, :16].permute(2, 3, 1, 0).reshape(49, 2, 2, 4, 2, 8)  # [tap][ks][h][t][nt][g]; cin = ks*8+h*4+t, cout = nt*8+g
    frag = wt.permute(0, 1, 5, 3, 4, 2).reshape(-1)                # [tap][ks][g][t][nt][h]
    if weight.shape[1] == 20:
        w4 = w[:, 16:].permute(2, 3, 1, 0).reshape(49, 4, 2, 8)    # [tap][t][nt][g]; cin = 16+t
        frag = torch.cat([frag, w4.permute(0, 3, 1, 2).reshape(-1)])   # [tap][g][t][nt]
    return frag.contiguous()


def conv7x7_c16_tf32(x: torch.Tensor, wfrag: torch.Tensor, bias: torch.Tensor, relu: bool = True) -> torch.Tensor:
    """conv 7x7 (16 or 20 -> 16 channels, pad 3) + bias (+ ReLU) on TF32 tensor cores; x and the result are
    channels-last [N,C,H,W]."""
    x = _nhwc(x, "activation")
    N, C, H, W = (int(v) for v in x.shape)
    if C not in (16, 20):
        raise L.FvpError(f"conv7x7_c16_tf32: 16 or 20 input channels expected, got {C}")
    wfrag, bias = L.require_cuda(wfrag, "filter fragments"), L.require_cuda(bias, "bias")
    if wfrag.numel() != 49 * 2 * 32 * 4 + (49 * 32 * 2 if C == 20 else 0) or bias.numel() != 16:
        raise L.FvpError("conv7x7_c16_tf32: bad filter fragment / bias size")
    out = torch.empty((N, H, W, 16), dtype=torch.float32, device=x.device).permute(0, 3, 1, 2)
    if out.numel():
        tok = TIMER.start("conv7x7_tf32", x.device)
        fn = L.load().fvp_conv7x7_c16_tf32 if C == 16 else L.load().fvp_conv7x7_c20_tf32
        L.check(fn(L.ptr(x), L.ptr(wfrag), L.ptr(bias), N, H, W, int(relu), L.ptr(out), L.stream_handle(x.device)),
                "fvp_conv7x7_tf32")
        TIMER.stop(tok)
        _launched()
    return out


def _tf32_round(w: torch.Tensor) -> torch.Tensor:
    """Round fp32 to TF32 (10-bit mantissa), ties away from zero, as cvt.rna.tf32.f32 does."""
    return ((w.contiguous().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)


def res2d_pack(c1, c2, skip):
    """Folded filters of one residual block -> the packed buffers of ``fvp_res2d_block_tf32``.
    ``c1`` / ``c2``: (weight [32,Cin,3,3], bias [32]) / ([32,32,3,3], [32]); ``skip``: 1x1 filters [32,Cin,1,1] or None."""
    def taps(w):
        cout, cin = int(w.shape[0]), int(w.shape[1])
        w = _tf32_round(w.detach().float())
        return w.permute(2, 3, 1, 0).reshape(-1, cin // 4, 4, cout).permute(0, 1, 3, 2).contiguous()   # [tap][cin/4][cout][4]
    w1, w2 = taps(c1[0]), taps(c2[0])
    ws = None if skip is None else taps(skip)[0].contiguous()
    return {"cin": int(c1[0].shape[1]), "w1": w1, "b1": c1[1].detach().float().contiguous(), "w2": w2,
            "b2": c2[1].detach().float().contiguous(), "ws": ws}


def res2d_supported(cin: int, cout: int, H: int, W: int, has_skip: bool) -> bool:
    if cout != 32 or cin not in (16, 32) or (cin != 32 and not has_skip):
        return False
    return bool(L.load().fvp_res2d_supported(int(cin), int(H), int(W)))


def res2d_block_tf32(x: torch.Tensor, packed) -> torch.Tensor:
    """One residual block (two 3x3 convolutions + skip + ReLUs) on the tcgen05 tensor cores; x and the result are
    channels-last [N,C,H,W], 16 or 32 channels in, 32 out."""
    x = _nhwc(x, "activation")
    N, C, H, W = (int(v) for v in x.shape)
    if C != packed["cin"]:
        raise L.FvpError(f"res2d_block_tf32: {packed['cin']} input channels expected, got {C}")
    out = torch.empty((N, H, W, 32), dtype=torch.float32, device=x.device).permute(0, 3, 1, 2)
    if out.numel():
        tok = TIMER.start("res2d_block_tf32", x.device)
        L.check(L.load().fvp_res2d_block_tf32(L.ptr(x), N, H, W, C, L.ptr(packed["w1"]), L.ptr(packed["b1"]), L.ptr(packed["w2"]),
                                              L.ptr(packed["b2"]), L.ptr(packed["ws"]), L.ptr(out), L.stream_handle(x.device)),
                "fvp_res2d_block_tf32")
        TIMER.stop(tok)
        _launched()
    return out


def res_tail_pack(c2_weight: torch.Tensor) -> torch.Tensor:
    """[32,32,3,3] filters -> the tap operand of ``fvp_res_tail_tf32``: [9][8][32][4] = W[cout][4p+i][ky][kx] at
    [3ky+kx][p][cout][i], TF32-rounded."""
    if tuple(c2_weight.shape) != (32, 32, 3, 3):
        raise L.FvpError(f"res_tail_pack: expected [32,32,3,3] filters, got {tuple(c2_weight.shape)}")
    w = _tf32_round(c2_weight.detach().float())
    return w.permute(2, 3, 1, 0).reshape(9, 8, 4, 32).permute(0, 1, 3, 2).contiguous()


def res_tail_supported(H: int, W: int, cout: int) -> bool:
    return bool(L.load().fvp_res_tail_supported(int(H), int(W), int(cout)))


def res_tail_tf32(h: torch.Tensor, xres: torch.Tensor, up: torch.Tensor, w2packed: torch.Tensor, b2_host: np.ndarray,
                  w_host: np.ndarray, b_host: np.ndarray) -> torch.Tensor:
    """The tail of P2PNet in one tcgen05 kernel: ``out = output_layer(relu(conv3x3(h) + b2 + xres) + shuffle(up))``.
    ``h``, ``xres`` channels-last [N,32,H,W]; ``up`` channels-last [N,128,H/2,W/2] (the un-shuffled last up-sampling);
    ``w2packed`` from ``res_tail_pack``; ``b2_host`` [32], ``w_host`` [Cout,32], ``b_host`` [Cout] numpy.  Planar [N,Cout,H,W]."""
    h, xres, up = _nhwc(h, "first convolution"), _nhwc(xres, "block input"), _nhwc(up, "upsampled")
    N, C, H, W = (int(v) for v in h.shape)
    cout = int(b_host.shape[0])
    if C != 32 or tuple(xres.shape) != (N, 32, H, W) or tuple(up.shape) != (N, 128, H // 2, W // 2):
        raise L.FvpError(f"res_tail_tf32: h {tuple(h.shape)}, xres {tuple(xres.shape)}, up {tuple(up.shape)} do not belong together")
    b2 = np.ascontiguousarray(b2_host, dtype=np.float32)
    wo, bo = np.ascontiguousarray(w_host, dtype=np.float32), np.ascontiguousarray(b_host, dtype=np.float32)
    if b2.shape != (32,) or wo.shape != (cout, 32) or w2packed.numel() != 9 * 8 * 32 * 4:
        raise L.FvpError(f"res_tail_tf32: b2 {b2.shape}, w {wo.shape}, packed filters {w2packed.numel()} do not match 32 channels")
    out = torch.empty((N, cout, H, W), dtype=torch.float32, device=h.device)
    if out.numel():
        tok = TIMER.start("res_tail_tf32", h.device)
        L.check(L.load().fvp_res_tail_tf32(L.ptr(h), L.ptr(xres), L.ptr(up), N, H, W, L.ptr(w2packed), b2.ctypes.data,
                                           wo.ctypes.data, bo.ctypes.data, cout, L.ptr(out), L.stream_handle(h.device)),
                "fvp_res_tail_tf32")
        TIMER.stop(tok)
        _launched()
    return out



def pose_match(pred: torch.Tensor, gt: torch.Tensor, vis: torch.Tensor, ngt: torch.Tensor):
    """pred [F,P,J,5] fp32, gt [F,G,J,3] fp64 (mm, padded), vis [F,G,J] fp32, ngt [F] int32 -> (err [F,P] fp64: mean distance
    of every valid pose to its closest ground-truth person over that person's visible joints, +inf where there is nothing to
    match; best [F,P] int32: index of that person or -1).  The device-side part of ``evaluation.evaluate_panoptic``."""
    pred = L.require_cuda(pred, "pred")
    gt = L.require_cuda(gt, "gt", torch.float64)
    vis = L.require_cuda(vis, "vis")
    ngt = L.require_cuda(ngt, "ngt", torch.int32)
    F, P, J = int(pred.shape[0]), int(pred.shape[1]), int(pred.shape[2])
    G = int(gt.shape[1])
    if pred.shape[3] != 5 or tuple(gt.shape) != (F, G, J, 3) or tuple(vis.shape) != (F, G, J) or tuple(ngt.shape) != (F,):
        raise L.FvpError(f"pose_match: inconsistent shapes {tuple(pred.shape)}, {tuple(gt.shape)}, {tuple(vis.shape)}, {tuple(ngt.shape)}")
    err = torch.empty((F, P), dtype=torch.float64, device=pred.device)
    best = torch.empty((F, P), dtype=torch.int32, device=pred.device)
    if F:
        L.check(L.load().fvp_pose_match(L.ptr(pred), L.ptr(gt), L.ptr(vis), L.ptr(ngt), F, P, J, G, L.ptr(err), L.ptr(best),
                                        L.stream_handle(pred.device)), "fvp_pose_match")
        _launched()
    return err, best


def conv7_pairs(weight: torch.Tensor) -> torch.Tensor:
    """[16,Cin,7,7] filters (Cin 16 or 20) -> the pair-tap operand of ``fvp_conv7x7_tc_tf32``: [28][Cin/4][32][4], pair-tap
    4*ky + pr holds the filters of tap (ky, 2pr) in rows 0..15 and of tap (ky, 2pr+1) in rows 16..31 (zeros beside kx = 6);
    TF32-rounded.  For 20 channels one plane of zeros ([32][4]) follows: the second half of the last k-step."""
    if tuple(weight.shape) not in ((16, 16, 7, 7), (16, 20, 7, 7)):
        raise L.FvpError(f"conv7_pairs: expected [16,16,7,7] or [16,20,7,7] filters, got {tuple(weight.shape)}")
    cin = int(weight.shape[1])
    w = _tf32_round(weight.detach().float())                      # [cout][cin][ky][kx]
    wp = torch.zeros((16, cin, 7, 8), dtype=torch.float32, device=w.device)
    wp[..., :7] = w
    t = wp.reshape(16, cin // 4, 4, 7, 4, 2)                      # [cout][p][i][ky][pr][half]
    out = t.permute(3, 4, 1, 5, 0, 2).reshape(-1)                 # [ky][pr][p][half*16+cout][i]
    if cin == 20:
        out = torch.cat([out, torch.zeros(32 * 4, dtype=torch.float32, device=w.device)])
    return out.contiguous()


def conv7x7_tc_supported(cin: int, H: int, W: int) -> bool:
    return bool(L.load().fvp_conv7x7_tc_supported(int(cin), int(H), int(W)))


def conv7x7_tc_tf32(x: torch.Tensor, wpairs: torch.Tensor, bias: torch.Tensor, relu: bool = True) -> torch.Tensor:
    """conv 7x7 (16 or 20 -> 16 channels, pad 3) + bias (+ ReLU) on the tcgen05 tensor cores; x and the result channels-last."""
    x = _nhwc(x, "activation")
    N, C, H, W = (int(v) for v in x.shape)
    if C not in (16, 20) or wpairs.numel() != 28 * (C // 4) * 32 * 4 + (128 if C == 20 else 0) or bias.numel() != 16:
        raise L.FvpError(f"conv7x7_tc_tf32: 16 / 20 channels with matching pair-taps and 16 biases expected (got {C}, {wpairs.numel()})")
    out = torch.empty((N, H, W, 16), dtype=torch.float32, device=x.device).permute(0, 3, 1, 2)
    if out.numel():
        tok = TIMER.start("conv7x7_tf32", x.device)
        L.check(L.load().fvp_conv7x7_tc_tf32(L.ptr(x), L.ptr(wpairs), L.ptr(bias), N, H, W, C, int(relu), L.ptr(out),
                                             L.stream_handle(x.device)), "fvp_conv7x7_tc_tf32")
        TIMER.stop(tok)
        _launched()
    return out
