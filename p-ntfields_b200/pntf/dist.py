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


def broadcast_replica_state(tensors, optimizer=None, src=0):
    """Make every rank a replica of rank `src`: weights, optimiser moments, Fourier matrices (in place) and the
    optimiser's step count.  Each rank initialises from its own random stream (Model.init_network), so without
    this the all-reduced gradient would belong to no single model."""
    rank, ws = world()
    if ws == 1:
        return
    for t in tensors:
        dist.broadcast(t, src)
    if optimizer is not None:
        c = torch.tensor([optimizer.step_count], dtype=torch.int64, device=tensors[0].device)
        dist.broadcast(c, src)
        optimizer.step_count = int(c.item())


class GradAllReduce:
    """All-reduce(sum) of grad[576129] and the loss scalars as ONE message.

    FlatParams keeps the gradient and the step's two loss scalars in one buffer (`gradbuf` = grad ++ loss), which
    the loss / reverse / weight-gradient kernels write directly: `reduce_flat` all-reduces that buffer in place, no
    packing copies.  `__call__` is the generic form for separate tensors (packs into an own buffer)."""

    def __init__(self, n_params, device, extra=2):
        self.n = n_params
        self.extra = extra
        self.device = device
        self.buf = None

    def reduce_flat(self, fp):
        """In place on fp.gradbuf; returns (grad, loss scalars averaged over the ranks)."""
        rank, ws = world()
        if ws == 1:
            return fp.grad, fp.loss_slot
        dist.all_reduce(fp.gradbuf, op=dist.ReduceOp.SUM)
        return fp.grad, fp.loss_slot / ws

    def __call__(self, grad, scalars):
        rank, ws = world()
        if ws == 1:
            return grad, scalars
        if self.buf is None:
            self.buf = torch.zeros(self.n + self.extra, dtype=torch.float32, device=self.device)
        self.buf[:self.n].copy_(grad)
        self.buf[self.n:].copy_(scalars)
        dist.all_reduce(self.buf, op=dist.ReduceOp.SUM)
        grad.copy_(self.buf[:self.n])
        return grad, self.buf[self.n:] / ws


def shard_rows(t, rank=None, world_size=None):
    """This rank's contiguous slice of the rows of t (queries of a sharded plan / field evaluation) and its bounds."""
    if rank is None:
        rank, world_size = world()
    b, e = shard_range(t.shape[0], rank, world_size)
    return t[b:e], b, e


def gather_rows(local, n_total, dim0_bounds=None):
    """All-gather row shards produced with shard_rows back into [n_total, ...] on every rank (optional: planning and
    inference need no exchange; callers that want the whole result on one rank ask for it)."""
    rank, ws = world()
    if ws == 1:
        return local
    counts = [shard_range(n_total, r, ws) for r in range(ws)]
    width = max(e - b for b, e in counts)
    pad = torch.zeros((width,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(ws)]
    dist.all_gather(parts, pad)
    return torch.cat([parts[r][:counts[r][1] - counts[r][0]] for r in range(ws)])
