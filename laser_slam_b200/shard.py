"""Multi-GPU plumbing of the scan-matching batch: scan pairs are independent, so a batch of B jobs is
cut into contiguous blocks, rank r owning [r*B/P, (r+1)*B/P) with its scans resident on its own GPU, and the only
exchange is an all-gather of the 128-byte result records for the pose graph.  torch.distributed is the transport
(NCCL on GPUs, gloo in the CPU tests); there is no data-path collective."""
import numpy as np

RECORD_BYTES = 128


def partition(total, world, rank):
    """Contiguous block of rank `rank`: [lo, hi).  Blocks differ by at most one job; multi-start callers pass the
    number of PAIRS so that all guesses of a pair stay on one GPU."""
    if world <= 0 or not (0 <= rank < world) or total < 0:
        raise ValueError("bad partition arguments")
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def block_sizes(total, world):
    return [partition(total, world, r)[1] - partition(total, world, r)[0] for r in range(world)]


def all_gather_records(local, total, group=None):
    """local: uint8 tensor [b, 128] of this rank's result records (b = its partition size).  Returns the
    [total, 128] tensor of all ranks' records in job order, on every rank.  Equal blocks use
    all_gather_into_tensor (one NCCL call); ragged blocks are padded to the largest block."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    if world == 1:
        return local
    sizes = block_sizes(total, world)
    assert local.shape[0] == sizes[dist.get_rank(group)] and local.shape[1] == RECORD_BYTES
    if len(set(sizes)) == 1:
        out = torch.empty((total, RECORD_BYTES), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    m = max(sizes)
    pad = torch.zeros((m, RECORD_BYTES), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)], 0)


def best_of_starts(records, starts):
    """Multi-start reduction (config 5): records = structured array [pairs*starts] in pair-major order; per pair keep
    the valid start with the smallest mean error (error / nvalid), ties to the first.  Returns (best records, index)."""
    r = records.reshape(-1, starts)
    score = np.where(r["valid"] == 1, r["error"] / np.maximum(r["nvalid"], 1), np.inf)
    idx = np.argmin(score, 1)
    return r[np.arange(r.shape[0]), idx], idx


class OverlappedGather:
    """The record all-gather of step k on a SIDE stream while step k+1 computes ("on a side stream,
    overlapped with the next batch's compute").  The caller keeps `depth` local record buffers and rotates them:

        k = g.slot()              # buffer this step writes; the compute stream first waits until the gather that last
                                  # read buffer k is done (an event, no host block)
        ... kernels write local[k] on the current stream ...
        g.submit(local[k])        # gather local[k] into out[k] on the side stream, behind an event of the current stream
        ...
        g.result(k)               # [total, 128] of every rank, in job order; the current stream waits for that gather

    Equal blocks only (the benchmark's weak-scaling shape).  On CPU tensors (gloo tests) there are no streams: submit
    gathers at once."""

    def __init__(self, total, like, depth=2, group=None):
        import torch
        import torch.distributed as dist
        self.group, self.depth, self.k = group, depth, -1
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.cuda = like.is_cuda
        self.out = [torch.empty((total, RECORD_BYTES), dtype=like.dtype, device=like.device) if self.world > 1 else None
                    for _ in range(depth)]
        self.done = [None] * depth
        self.side = torch.cuda.Stream(device=like.device) if (self.cuda and self.world > 1) else None

    def slot(self):
        import torch
        self.k = (self.k + 1) % self.depth
        if self.done[self.k] is not None:
            torch.cuda.current_stream().wait_event(self.done[self.k])
        return self.k

    def submit(self, local):
        import torch
        import torch.distributed as dist
        k = self.k
        if self.world == 1:
            self.out[k] = local
            return k
        assert local.shape[0] * self.world == self.out[k].shape[0] and local.shape[1] == RECORD_BYTES
        if self.side is None:
            dist.all_gather_into_tensor(self.out[k], local.contiguous(), group=self.group)
            return k
        ready = torch.cuda.Event()
        ready.record()                      # the records of this step are complete at this point of the compute stream
        with torch.cuda.stream(self.side):
            self.side.wait_event(ready)
            dist.all_gather_into_tensor(self.out[k], local, group=self.group)
            ev = torch.cuda.Event()
            ev.record(self.side)
        self.done[k] = ev
        return k

    def result(self, k):
        import torch
        if self.side is not None and self.done[k] is not None:
            torch.cuda.current_stream().wait_event(self.done[k])
        return self.out[k]
