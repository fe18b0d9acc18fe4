"""CPU tests of the multi-GPU host logic (beifong_b200/multi.py) over the gloo backend, world size 2:
sample-index sharding + one reduce equals the unsharded profile.  The per-rank compute here is the CPU oracle
(test infrastructure); the GPU path runs the same host logic around bf_receive (bench.py --gpus N)."""
import os
import socket

import numpy as np
import pytest

from beifong_b200 import multi


def test_shard_samples_partition():
    for n in (0, 1, 7, 8, 1000, 4096 * 3 + 1):
        for world in (1, 2, 3, 8):
            parts = [multi.shard_samples(n, r, world) for r in range(world)]
            assert parts[0][0] == 0 and sum(c for _, c in parts) == n
            for (f0, c0), (f1, _) in zip(parts, parts[1:]):
                assert f1 == f0 + c0
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
    with pytest.raises(ValueError):
        multi.shard_samples(10, 2, 2)


def test_shard_items_round_robin():
    got = sorted(sum((multi.shard_items(256, r, 8) for r in range(8)), []))
    assert got == list(range(256))
    assert multi.shard_items(3, 5, 8) == []


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, n_total, seed, out_dir):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here); sys.path.insert(0, os.path.dirname(here))
    import torch
    import torch.distributed as dist
    from beifong_b200 import configs, multi
    from oracle_lib import OracleScene
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = OracleScene(configs.small_radar().export(), accel=OracleScene.BVH)
    first, n = multi.shard_samples(n_total, rank, world)
    prof = torch.from_numpy(orc.receive(n, seed=seed, first=first, threads=1).copy())
    is_dst = multi.reduce_profile(prof, dst=0)
    if is_dst:
        np.save(os.path.join(out_dir, "reduced.npy"), prof.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_reduce_matches_single(tmp_path):
    import torch.multiprocessing as mp
    from beifong_b200 import configs
    from oracle_lib import OracleScene
    n_total, seed, world = 30001, 4, 2
    mp.spawn(_worker, args=(world, _free_port(), n_total, seed, str(tmp_path)), nprocs=world, join=True)
    reduced = np.load(tmp_path / "reduced.npy")
    whole = OracleScene(configs.small_radar().export(), accel=OracleScene.BVH).receive(n_total, seed=seed, threads=1)
    assert reduced.shape == whole.shape
    assert np.array_equal(reduced[..., 2], whole[..., 2]) and np.array_equal(reduced[..., 1], whole[..., 1])
    # fp32 summation order differs between the sharded and the single run: <= 1e-3 relative L2 (north_star), not bitwise
    err = np.linalg.norm(reduced[..., 0].astype(np.float64) - whole[..., 0]) / np.linalg.norm(whole[..., 0].astype(np.float64))
    assert err <= 1e-5


class _OracleFrames:
    """The slice of FlatScene that multi.receive_sharded drives (receive with accumulate / clear), computed by the CPU oracle
    into a torch tensor that stands for the device profile."""
    def __init__(self, orc, shape):
        import torch
        self.orc, self.profile = orc, torch.zeros(shape, dtype=torch.float32)

    def receive(self, n, seed=0, first_sample=0, accumulate=False):
        import torch
        if not accumulate:
            self.profile.zero_()
        if n > 0:
            self.profile += torch.from_numpy(self.orc.receive(n, seed=seed, first=first_sample, threads=1).copy())


def _worker_accumulate(rank, world, port, frames, out_dir):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here); sys.path.insert(0, os.path.dirname(here))
    import torch.distributed as dist
    from beifong_b200 import configs, multi
    from oracle_lib import OracleScene
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = OracleScene(configs.small_radar().export(), accel=OracleScene.BVH)
    shape = orc.receive(1, seed=0, threads=1).shape
    scene = _OracleFrames(orc, shape)
    for k, (n_total, seed) in enumerate(frames):
        multi.receive_sharded(scene, n_total, seed, rank, world, accumulate=k > 0, profile_tensor=scene.profile, stream=None, dst=0)
    if rank == 0:
        np.save(os.path.join(out_dir, "accumulated.npy"), scene.profile.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_accumulated_frames_match_single(tmp_path):
    """Two sharded frames accumulated on rank 0 (receive_sharded(accumulate=True): the other rank clears its partial sum after
    each in-place reduce) equal the two frames accumulated by one unsharded run -- the double count ADVICE r01 found."""
    import torch.multiprocessing as mp
    from beifong_b200 import configs
    from oracle_lib import OracleScene
    frames, world = [(20001, 4), (15000, 9)], 2
    mp.spawn(_worker_accumulate, args=(world, _free_port(), frames, str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / "accumulated.npy")
    orc = OracleScene(configs.small_radar().export(), accel=OracleScene.BVH)
    whole = sum(orc.receive(n, seed=seed, threads=1).astype(np.float64) for n, seed in frames)
    assert np.array_equal(got[..., 2], whole[..., 2]) and np.array_equal(got[..., 1], whole[..., 1])
    err = np.linalg.norm(got[..., 0].astype(np.float64) - whole[..., 0]) / np.linalg.norm(whole[..., 0])
    assert err <= 1e-5
