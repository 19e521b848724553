"""Host logic of the multi-GPU partitioning on CPU: world-size-2 (and 3) gloo process groups."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from faster_voxelpose_b200 import parallel as par


def test_partitions_cover_exactly_once():
    for n in (0, 1, 5, 20, 64, 97):
        for world in (1, 2, 3, 4, 8):
            got = []
            for r in range(world):
                lo, hi = par.zslab(n, r, world)
                assert 0 <= lo <= hi <= n
                got += list(range(lo, hi))
                s = par.frame_shard(n, r, world)
                assert (s.start, s.stop) == (lo, hi)
            assert got == list(range(n))
            sizes = [par.zslab(n, r, world)[1] - par.zslab(n, r, world)[0] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        par.zslab(10, 2, 2)


def test_peer_plane_layout_is_rank_independent_and_aligned():
    """Staging layout of PeerPlanes: offsets identical on every rank (they index peers' buffers), 16-byte
    aligned regions, slabs sized for the longest slab; zslab_starts are the boundaries zslab() hands out."""
    for Z, world in ((64, 2), (64, 8), (20, 3), (5, 8)):
        starts = par.zslab_starts(Z, world)
        assert starts[0] == 0 and starts[-1] == Z and len(starts) == world + 1
        assert [(starts[r], starts[r + 1]) for r in range(world)] == [par.zslab(Z, r, world) for r in range(world)]
        offs, total = par.peer_plane_layout(3, 15, 7, 9, Z, world)
        longest = max(starts[r + 1] - starts[r] for r in range(world))
        assert all(o % 4 == 0 for o in offs) and total % 4 == 0
        assert offs[0] == 0 and offs[1] >= 3 * 15 * 7 * 9 and offs[2] - offs[1] >= 3 * 15 * 7 * longest
        assert offs[3] - offs[2] >= 3 * 15 * 9 * longest
        assert offs[4] - offs[3] >= 3 * 15 * 7 * 9 and offs[5] - offs[4] >= 3 * 15 * 7 * Z and total - offs[5] >= 3 * 15 * 9 * Z
        offs_xy, total_xy = par.peer_plane_layout(3, 15, 7, 9, Z, world, want=(True, False, False))
        assert total_xy == 2 * ((3 * 15 * 7 * 9 + 3) // 4 * 4) and offs_xy[1] == offs_xy[2] == offs_xy[3]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, nz, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        gen = torch.Generator().manual_seed(1234)                  # same cube on every rank
        cube = torch.rand((2, 3, 6, 5, nz), generator=gen)
        z0, z1 = par.zslab(nz, rank, world)
        slab = cube[..., z0:z1]
        xy = slab.max(dim=4)[0] if z1 > z0 else torch.zeros(cube.shape[:4])
        xz, yz = slab.max(dim=3)[0], slab.max(dim=2)[0]            # [2,3,6,zs], [2,3,5,zs]
        fxy, fxz, fyz = par.combine_zslab_planes(xy.clone(), xz, yz, nz)
        ok = torch.equal(fxy, cube.max(dim=4)[0]) and torch.equal(fxz, cube.max(dim=3)[0]) \
            and torch.equal(fyz, cube.max(dim=2)[0])
        # frame sharding + pose gather
        frames = 5
        poses = torch.arange(frames * 2 * 3 * 5, dtype=torch.float32).view(frames, 2, 3, 5)
        mine = poses[par.frame_shard(frames, rank, world)]
        ok = ok and torch.equal(par.gather_poses(mine, frames), poses)
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,nz", [(2, 20), (2, 7), (3, 8), (3, 2)])
def test_zslab_planes_and_pose_gather_over_gloo(world, nz):
    port = _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, port, nz, ret), nprocs=world, join=True)
        assert dict(ret) == {r: True for r in range(world)}
