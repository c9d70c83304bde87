"""Data-parallel plumbing: one process per GPU, torch.distributed (NCCL over NVLink; gloo in CPU
tests).  The only exchange on the hot path is one all-reduce(sum) per step over the flat gradient
buffer with the two loss scalars appended; planning and inference shard
queries with no communication."""
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n, rank, world_size):
    """Contiguous, balanced [begin, end) slice of n items for this rank."""
    base, rem = divmod(n, world_size)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


class GradAllReduce:
    """Packs grad[576129] + loss scalars into one buffer and all-reduces it (sum)."""

    def __init__(self, n_params, device, extra=2):
        self.buf = torch.zeros(n_params + extra, dtype=torch.float32, device=device)
        self.n = n_params

    def __call__(self, grad, scalars):
        rank, ws = world()
        if ws == 1:
            return grad, scalars
        self.buf[:self.n].copy_(grad)
        self.buf[self.n:].copy_(scalars)
        dist.all_reduce(self.buf, op=dist.ReduceOp.SUM)
        grad.copy_(self.buf[:self.n])
        return grad, self.buf[self.n:] / ws
