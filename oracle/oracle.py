"""TEST INFRASTRUCTURE -- ctypes/numpy front end of the C oracle (``fvp_oracle.c``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` leg import this.
The product (``faster_voxelpose_b200``) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(_HERE, "fvp_oracle.c")
LIB = os.path.join(_HERE, "_build", "libfvp_oracle.so")
CAM_STRIDE = 24


def build(force: bool = False) -> str:
    """gcc the restatement into oracle/_build/ (git-ignored; travels to the GPU box)."""
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.check_call(["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC",
                           "-o", LIB, SRC, "-lm"])
    return LIB


class Geom(C.Structure):
    _fields_ = [("V", C.c_int), ("J", C.c_int), ("H", C.c_int), ("W", C.c_int), ("clamp_max", C.c_float),
                ("t", C.c_float * 6), ("hm_w", C.c_float), ("hm_h", C.c_float), ("img_w", C.c_float),
                ("img_h", C.c_float), ("wm1", C.c_float), ("hm1", C.c_float)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        _lib = C.CDLL(LIB)
    return _lib


def _p(a, t=C.c_float):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def pack_cameras(cams: Sequence[Dict]) -> np.ndarray:
    """[V,24] fp32 camera table: R9 T3 fx fy cx cy k3 p2 pad3 (float64 -> fp32 casts exactly like
    unfold_camera_param, cameras.py:11-18)."""
    out = np.zeros((len(cams), CAM_STRIDE), dtype=np.float32)
    for v, c in enumerate(cams):
        out[v, 0:9] = np.asarray(c["R"], dtype=np.float64).reshape(9).astype(np.float32)
        out[v, 9:12] = np.asarray(c["T"], dtype=np.float64).reshape(3).astype(np.float32)
        out[v, 12:16] = np.array([c["fx"], c["fy"], c["cx"], c["cy"]], dtype=np.float64).astype(np.float32)
        out[v, 16:19] = np.asarray(c["k"], dtype=np.float64).reshape(3).astype(np.float32)
        out[v, 19:21] = np.asarray(c["p"], dtype=np.float64).reshape(2).astype(np.float32)
    return out


def make_geom(cfg, V: int, resize_transform) -> Geom:
    ds = cfg.DATASET
    w, h = ds.HEATMAP_SIZE
    g = Geom()
    g.V, g.J, g.H, g.W = V, ds.NUM_JOINTS, h, w
    g.clamp_max = float(max(ds.ORI_IMAGE_SIZE[0], ds.ORI_IMAGE_SIZE[1]))
    t = np.asarray(resize_transform, dtype=np.float64).astype(np.float32).reshape(6)
    for i in range(6):
        g.t[i] = float(t[i])
    g.hm_w, g.hm_h = float(w), float(h)
    g.img_w, g.img_h = float(ds.IMAGE_SIZE[0]), float(ds.IMAGE_SIZE[1])
    g.wm1, g.hm1 = float(w - 1), float(h - 1)
    return g


def linspace_f32(lo: float, hi: float, n: int) -> np.ndarray:
    """torch.linspace(lo, hi, n) in fp32.  ATen's CPU kernel mixes a vectorised ``arange`` with
    FMA-contracted scalar tails, so its low bits depend on the host's vector ISA (AVX2 vs
    AVX-512 differ); it is not restated -- torch itself is called, exactly like the reference
    does (project_whole.py:62-64).  Golden comparisons use the tables stored in
    tests/golden/*.npz instead (``whole_axes`` / ``fine_axes`` / ``ind_axes``)."""
    import torch

    return torch.linspace(lo, hi, int(n)).numpy().astype(np.float32)


def axis_tables(space_size, space_center, voxels):
    """project_whole.py:62-70 -- three 1-D tables of voxel-centre coordinates."""
    return [(linspace_f32(-space_size[a] / 2, space_size[a] / 2, int(voxels[a]))
             + np.float32(space_center[a])).astype(np.float32) for a in range(3)]


class IndividualSpec:
    """project_individual.py:43-111 constants in numpy fp32 / int32."""

    def __init__(self, cfg):
        f32 = np.float32
        self.whole_center = np.asarray(cfg.CAPTURE_SPEC.SPACE_CENTER, dtype=f32)
        self.whole_size = np.asarray(cfg.CAPTURE_SPEC.SPACE_SIZE, dtype=f32)
        self.ind_size = np.asarray(cfg.INDIVIDUAL_SPEC.SPACE_SIZE, dtype=f32)
        self.voxels = np.asarray(cfg.INDIVIDUAL_SPEC.VOXELS_PER_AXIS, dtype=np.int32)
        self.fine = ((self.whole_size / self.ind_size) * (self.voxels - 1).astype(f32)).astype(np.int32) + 1
        self.scale = ((self.fine.astype(f32) - f32(1)) / self.whole_size).astype(f32)
        self.bias = (-(self.ind_size / f32(2.0)) / self.whole_size * (self.fine - 1).astype(f32)
                     - self.scale * (self.whole_center - self.whole_size / f32(2.0))).astype(f32)
        self.fine_axes = axis_tables(self.whole_size.tolist(), self.whole_center.tolist(), self.fine)
        ax, ay, az = axis_tables(self.ind_size.tolist(), self.whole_center.tolist(), self.voxels)
        n = [int(v) for v in self.voxels]
        xy = np.stack(np.meshgrid(ax, ay, indexing="ij"), axis=-1).reshape(-1, 2)
        xz = np.stack(np.meshgrid(ax, az, indexing="ij"), axis=-1).reshape(-1, 2)
        yz = np.stack(np.meshgrid(ay, az, indexing="ij"), axis=-1).reshape(-1, 2)
        assert n[0] == n[1] == n[2], "center_grid stacking needs equal voxel counts (as in the reference)"
        self.center_grid = np.stack([xy, xz, yz]).astype(f32)


def sample_grid(cams_packed, geom: Geom, axes, dims) -> np.ndarray:
    X, Y, Z = [int(d) for d in dims]
    out = np.empty((geom.V, X * Y * Z, 2), dtype=np.float32)
    ax, ay, az = [_f32(a) for a in axes]
    lib().orc_sample_grid(_p(_f32(cams_packed)), C.byref(geom), _p(ax), _p(ay), _p(az), X, Y, Z, _p(out))
    return out


def project_cubes(heatmaps, rigs_packed, frame_rig, geom: Geom, axes, frame, tl, start, end, skip, dims):
    nx, ny, nz = [int(d) for d in dims]
    hm = _f32(heatmaps)
    frame, tl, start, end = _i32(frame), _i32(tl), _i32(start), _i32(end)
    skip = _i32(skip) if skip is not None else None
    ncube = frame.shape[0]
    out = np.empty((ncube, geom.J, nx, ny, nz), dtype=np.float32)
    ax, ay, az = [_f32(a) for a in axes]
    lib().orc_project_cubes(_p(hm), _p(_f32(rigs_packed)), _p(_i32(frame_rig), C.c_int), C.byref(geom), _p(ax), _p(ay),
                            _p(az), ncube, _p(frame, C.c_int), _p(tl, C.c_int), _p(start, C.c_int),
                            _p(end, C.c_int), _p(skip, C.c_int), nx, ny, nz, _p(out))
    return out


def project_whole(heatmaps, rigs_packed, frame_rig, geom: Geom, axes, dims) -> np.ndarray:
    """a5: space cube [B,J,X,Y,Z]."""
    B = heatmaps.shape[0]
    z3 = np.zeros((B, 3), dtype=np.int32)
    end = np.tile(np.asarray(dims, dtype=np.int32), (B, 1))
    return project_cubes(heatmaps, rigs_packed, frame_rig, geom, axes, np.arange(B), z3, z3, end, None, dims)


def individual_windows(spec: IndividualSpec, props: np.ndarray):
    props = _f32(props).reshape(-1, 7)
    n = props.shape[0]
    tl = np.empty((n, 3), np.int32)
    start = np.empty((n, 3), np.int32)
    end = np.empty((n, 3), np.int32)
    skip = np.empty((n,), np.int32)
    offset = np.empty((n, 3), np.float32)
    lib().orc_individual_windows(_p(props), n, _p(spec.scale), _p(spec.bias), _p(_i32(spec.fine), C.c_int),
                                 _p(spec.whole_size), _p(spec.ind_size), _p(_i32(spec.voxels), C.c_int),
                                 _p(tl, C.c_int), _p(offset), _p(start, C.c_int), _p(end, C.c_int), _p(skip, C.c_int))
    return tl, offset, start, end, skip


def project_individual(heatmaps, rigs_packed, frame_rig, geom: Geom, spec: IndividualSpec, props, frame):
    """a12: cubes [N,J,nx,ny,nz], offset [N,3] for N proposals; frame[n] = frame index."""
    tl, offset, start, end, skip = individual_windows(spec, props)
    cubes = project_cubes(heatmaps, rigs_packed, frame_rig, geom, spec.fine_axes, frame, tl, start, end, skip,
                          spec.voxels)
    return cubes, offset


def planes3(cubes: np.ndarray):
    cubes = _f32(cubes)
    n, j, nx, ny, nz = cubes.shape
    xy = np.empty((n, j, nx, ny), np.float32)
    xz = np.empty((n, j, nx, nz), np.float32)
    yz = np.empty((n, j, ny, nz), np.float32)
    lib().orc_planes3(_p(cubes), n * j, nx, ny, nz, _p(xy), _p(xz), _p(yz))
    return xy, xz, yz


def nms_topk(hm: np.ndarray, P: int):
    hm = _f32(hm)
    B, X, Y = hm.shape[0], hm.shape[-2], hm.shape[-1]
    vals = np.empty((B, P), np.float32)
    idx2 = np.empty((B, P, 2), np.int64)
    flat = np.empty((B, P), np.int64)
    lib().orc_nms_topk(_p(hm), B, X, Y, P, _p(vals), _p(idx2, C.c_int64), _p(flat, C.c_int64))
    return vals, idx2, flat


def proposal_records(idx2, conf2d, hm1d, bbox, cfg) -> np.ndarray:
    cs = cfg.CAPTURE_SPEC
    f32 = np.float32
    scale = (np.asarray(cs.SPACE_SIZE, f32) / (np.asarray(cs.VOXELS_PER_AXIS) - 1).astype(f32)).astype(f32)
    bias = (np.asarray(cs.SPACE_CENTER, f32) - np.asarray(cs.SPACE_SIZE, f32) / f32(2.0)).astype(f32)
    idx2 = np.ascontiguousarray(idx2, dtype=np.int64)
    B, P = conf2d.shape
    hm1d = _f32(hm1d)
    rec = np.empty((B, P, 7), f32)
    lib().orc_proposal_records(_p(idx2, C.c_int64), _p(_f32(conf2d)), _p(hm1d), _p(_f32(bbox)), B * P, hm1d.shape[-1],
                               _p(scale), _p(bias), C.c_float(cs.MIN_SCORE), _p(rec))
    return rec


def softargmax(feat: np.ndarray, center_grid: np.ndarray, beta: float):
    feat = _f32(feat)
    _, N, J = feat.shape[:3]
    M = int(np.prod(feat.shape[3:]))
    pose = np.empty((3, N, J, 2), np.float32)
    conf = np.empty((N,), np.float32)
    lib().orc_softargmax(_p(feat), _p(_f32(center_grid)), N, J, M, C.c_float(beta), _p(pose), _p(conf))
    return pose, conf


def fuse(pose: np.ndarray, offset: np.ndarray, weights: np.ndarray):
    """Returns (pose_with_offsets [3,N,J,2], fused [N,J,3])."""
    pose = _f32(pose).copy()
    _, N, J = pose.shape[:3]
    fused = np.empty((N, J, 3), np.float32)
    lib().orc_fuse(_p(pose), _p(_f32(offset)), _p(_f32(weights).reshape(3, N, J)), N, J, _p(fused))
    return pose, fused
