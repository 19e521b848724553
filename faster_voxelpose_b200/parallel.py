"""Multi-GPU partitioning of the hot path: one process per GPU, torch.distributed.

* Frames are independent (the reference loops over them, project_whole.py:142,
  joint_localization_net.py:148): ``frame_shard`` gives each rank its slice of a batch; no data-path
  collective is needed, only the result gather (``gather_poses``).
* Dense space grids are split into z slabs: rank r projects z in ``zslab(nz, r, world)`` with
  ``fvp_project_planes`` and owns a partial xy max-plane plus its slab of the xz / yz planes;
  ``combine_zslab_planes`` turns them into the full planes on every rank with one
  ``all_reduce(MAX)`` (xy) and one ``all_gather`` (xz and yz slabs, padded to equal length).

The collectives go through whatever backend the process group was created with: NCCL over
NVLink/NVSwitch on the GPU box, gloo in the CPU tests of this host logic.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch
import torch.distributed as dist

__all__ = ["frame_shard", "zslab", "combine_zslab_planes", "gather_poses", "project_space_planes_zsharded"]


def _split(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block partition of range(n): the first n % world ranks get one extra element."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(int(n), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def frame_shard(num_frames: int, rank: int, world: int) -> slice:
    """Frames of a global batch owned by ``rank`` (may be empty when num_frames < world)."""
    lo, hi = _split(num_frames, rank, world)
    return slice(lo, hi)


def zslab(nz: int, rank: int, world: int) -> Tuple[int, int]:
    """z range [z0, z1) of the space grid owned by ``rank`` (empty slabs when nz < world)."""
    return _split(nz, rank, world)


def _world(group) -> Tuple[int, int]:
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def combine_zslab_planes(xy_partial: torch.Tensor, xz_slab: Optional[torch.Tensor], yz_slab: Optional[torch.Tensor],
                         nz: int, group=None, z_dim: int = -1):
    """Partial planes of this rank's z slab -> full planes on every rank.

    ``xy_partial`` [..] is the max over the slab: combined in place with all_reduce(MAX) (values are
    >= 0 or NaN; an empty slab contributes zeros).  ``xz_slab`` / ``yz_slab`` hold the slab along
    ``z_dim`` (length z1 - z0, possibly 0): they are padded to the longest slab, all-gathered in one
    call each and trimmed back, giving tensors of length ``nz`` along ``z_dim``."""
    rank, world = _world(group)
    if world == 1:
        return xy_partial, xz_slab, yz_slab
    dist.all_reduce(xy_partial, op=dist.ReduceOp.MAX, group=group)
    longest = -(-int(nz) // world)

    def gather(slab):
        if slab is None:
            return None
        zd = z_dim % slab.dim()
        moved = slab.movedim(zd, 0)                                   # slab axis first: [zs, ...]
        padded = torch.zeros((longest,) + tuple(moved.shape[1:]), dtype=slab.dtype, device=slab.device)
        padded[:moved.shape[0]] = moved
        out = torch.empty((world * longest,) + tuple(moved.shape[1:]), dtype=slab.dtype, device=slab.device)
        dist.all_gather_into_tensor(out, padded.contiguous(), group=group)
        pieces = []
        for r in range(world):
            lo, hi = zslab(nz, r, world)
            pieces.append(out[r * longest:r * longest + (hi - lo)])
        return torch.cat(pieces, dim=0).movedim(0, zd).contiguous()

    return xy_partial, gather(xz_slab), gather(yz_slab)


def gather_poses(local_poses: torch.Tensor, num_frames: int, group=None) -> torch.Tensor:
    """[B_local,P,J,5] of every rank's frame shard -> [num_frames,P,J,5] on every rank (30 KB/frame)."""
    rank, world = _world(group)
    if world == 1:
        return local_poses
    longest = -(-int(num_frames) // world)
    padded = torch.zeros((longest,) + tuple(local_poses.shape[1:]), dtype=local_poses.dtype, device=local_poses.device)
    padded[:local_poses.shape[0]] = local_poses
    out = torch.empty((world * longest,) + tuple(local_poses.shape[1:]), dtype=local_poses.dtype,
                      device=local_poses.device)
    dist.all_gather_into_tensor(out, padded, group=group)
    pieces = []
    for r in range(world):
        s = frame_shard(num_frames, r, world)
        pieces.append(out[r * longest:r * longest + (s.stop - s.start)])
    return torch.cat(pieces, dim=0)


def project_space_planes_zsharded(layer, hm_cl, frame_rig, rigs, geom, group=None, want=(True, True, True)):
    """z-sharded fused projection of the space grid (BASELINE config 4): every rank holds the same
    frames (heatmaps replicated), projects its z slab with the fused kernel and the planes are
    combined over the group.  Returns full (xy [B,J,X,Y], xz [B,J,X,Z], yz [B,J,Y,Z])."""
    from . import ops

    rank, world = _world(group)
    X, Y, Z = [int(v) for v in layer.voxels_per_axis]
    z0, z1 = zslab(Z, rank, world)
    dev = hm_cl.device
    B, J = hm_cl.shape[0], geom.J
    if z1 > z0:
        grid = layer.cached_grids(rigs, geom, dev)
        xy, xz, yz = ops.project_planes(hm_cl, rigs.dev, frame_rig, geom, layer.axes(dev), [X, Y, Z], B,
                                        cached_grid=grid, mode=layer.mode, want=want, zslab=(z0, z1))
    else:
        xy = torch.zeros((B, J, X, Y), device=dev) if want[0] else None
        xz = torch.zeros((B, J, X, 0), device=dev) if want[1] else None
        yz = torch.zeros((B, J, Y, 0), device=dev) if want[2] else None
    if xy is None:
        raise ValueError("the xy plane is what the detection net consumes; want[0] must be True")
    return combine_zslab_planes(xy, xz, yz, Z, group=group)
