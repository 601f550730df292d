"""Host-side sharding of the association path over one process per GPU.

The cloud and its index are replicated; SfM points (queries) are split into contiguous shards, images
round-robin.  The path has exactly one exchange step per pass: the per-match [r, Jx, Jy, Jz] blocks are
gathered and sum r^2 / the accepted and ground counters are all-reduced.  Everything here works on any
torch.distributed backend (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def shard_bounds(n, world, rank):
    """Contiguous, balanced split of n units: the first n % world shards get one extra."""
    base, extra = divmod(int(n), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_images(n_img, world, rank):
    """Round-robin image assignment for the depth projection (450 images -> 57/56 per GPU at world 8)."""
    return list(range(rank, int(n_img), int(world)))


def reduce_stats(sum_sq, accepted, ground, group=None):
    """All-reduce (sum, fp64) of the residual norm and the two counters.  Returns a float64 tensor [3]."""
    t = sum_sq if torch.is_tensor(sum_sq) and sum_sq.numel() == 3 else None
    if t is None:
        dev = sum_sq.device if torch.is_tensor(sum_sq) else "cpu"
        t = torch.tensor([float(sum_sq), float(accepted), float(ground)], dtype=torch.float64, device=dev)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


class BlockGather:
    """Gathers equal-sized per-rank [Q, 4] fp64 blocks (r | J) into one [world * Q, 4] tensor on every rank.
    On CUDA the collective runs on its own stream so that it overlaps the next pass's kernels; wait()
    makes the caller's stream depend on it."""

    def __init__(self, q_per_rank, device, group=None):
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.group = group
        self.out = torch.empty((self.world * q_per_rank, 4), dtype=torch.float64, device=device)
        self.cuda = torch.device(device).type == "cuda"
        self.stream = torch.cuda.Stream(device=device) if self.cuda else None
        self.work = None

    def start(self, block):
        if self.world == 1:
            self.out.copy_(block)
            return
        if self.cuda:
            self.stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.stream):
                dist.all_gather_into_tensor(self.out, block, group=self.group)
        else:
            self.work = dist.all_gather_into_tensor(self.out, block, group=self.group, async_op=True)

    def wait(self):
        if self.world == 1:
            return self.out
        if self.cuda:
            torch.cuda.current_stream().wait_stream(self.stream)
        elif self.work is not None:
            self.work.wait(); self.work = None
        return self.out


def gather_ragged(local, group=None):
    """All-gather of per-rank 1-D/2-D tensors whose first dimension differs (result sets of a ragged shard)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local
    n = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    m = max(sizes)
    pad = torch.zeros((m,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)], 0)
