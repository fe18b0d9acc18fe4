"""Multi-GPU host logic: one process per GPU, scene replicated, sample indices sharded,
per-receiver profiles combined with ONE reduce (NCCL over NVLink on GPUs; the same code runs over gloo on
CPU tensors in the tests).  The reference has no counterpart: it is single-process.

Samples are independent given (seed, sample index) -- per-sample PCG32 streams, sampler.cpp:89-92 -- so any
partition of the index range [0, n_total) sums to the single-GPU result up to fp32 summation order.
"""
import numpy as np


def shard_samples(n_total, rank, world):
    """Contiguous, balanced partition of [0, n_total): returns (first_sample, n_samples) of `rank`.
    The first (n_total % world) ranks take one extra sample; empty shards are allowed (n_total < world)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"shard_samples: rank {rank} outside world of {world}")
    if n_total < 0:
        raise ValueError("shard_samples: negative sample count")
    base, extra = divmod(int(n_total), int(world))
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def shard_items(n_items, rank, world):
    """Round-robin assignment of independent units (frames of config C4, receiver elements of C5) to ranks."""
    return list(range(rank, n_items, world))


def reduce_profile(profile, dst=0, group=None):
    """Sum the per-rank range profiles into rank `dst` (torch tensor, in place).  A no-op without an initialised
    process group / at world size 1.  Returns True on the rank that holds the result."""
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return True
    dist.reduce(profile, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return dist.get_rank(group) == dst


class DeviceProfile:
    """__cuda_array_interface__ view of a committed scene's device profile, so that torch / NCCL can reduce it in
    place on the scene's own stream (no copy)."""

    def __init__(self, scene):
        ptr, n, stream = scene.profile_device()
        self.stream_ptr = stream
        self.n = n
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}

    def tensor(self, device):
        import torch
        return torch.as_tensor(self, device=device)

    def stream(self, device):
        import torch
        return torch.cuda.ExternalStream(self.stream_ptr, device=device)


def init_native_reduce(scene, rank, world):
    """Create the library's own NCCL communicator for `scene` (bf_multi_init).  The 128-byte unique id travels through
    the already initialised torch.distributed group (host-side plumbing only); the reduce itself is bf_multi_reduce."""
    import torch.distributed as dist
    box = [scene.multi_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    scene.multi_init(rank, world, box[0])


def receive_sharded(scene, n_total, seed, rank, world, accumulate=False, profile_tensor=None, stream=None, dst=0, native=False):
    """One sharded frame: this rank's slice of the sample range through bf_receive, then the single reduce, issued
    on the scene's stream so that it follows the last wavefront without a host sync.
    native=True: the reduce is bf_multi_reduce (after init_native_reduce); otherwise torch.distributed on profile_tensor.
    accumulate=True: rank `dst` keeps the running total.  The reduce is in place, so every other rank clears its partial sum
    after it has been reduced, on every frame (otherwise a later accumulated frame would add it a second time)."""
    first, n = shard_samples(n_total, rank, world)
    if n > 0:
        scene.receive(n, seed=seed, first_sample=first, accumulate=accumulate)
    elif not accumulate:
        scene.receive(0, seed=seed, first_sample=first, accumulate=False)
    if world > 1:
        if native:
            scene.multi_reduce(dst)
        elif profile_tensor is not None:
            import torch
            with torch.cuda.stream(stream):
                reduce_profile(profile_tensor, dst=dst)
        if rank != dst and (native or profile_tensor is not None):
            # clear, stream-ordered behind the reduce -- after EVERY frame, also one received with accumulate=False: the next
            # frame may accumulate, and this rank's partial sum is already part of dst's total
            scene.receive(0, seed=seed, first_sample=first, accumulate=False)
    return first, n
