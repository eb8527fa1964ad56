"""1-D row-block sharding of the pull products over the GPUs of one node.

Every rank owns a contiguous block of the OUTPUT rows of `w = P (+).(x) u` (rows of the pull matrix P, i.e. of A'
for `A.T.mxv` / `u.vxm(A)`) and the matching block of every vector; one all-gather per product rebuilds the
full-length input vector.  This module holds the host-side logic only (partition + exchange on torch tensors), so
it runs unchanged over NCCL on device buffers (bench.py) and over gloo on CPU tensors (tests/test_sharding_gloo.py).
"""
import torch
import torch.distributed as dist


def block_size(n, world):
    """Rows per rank: equal blocks of ceil(n / world) rows; the last ranks may own fewer (or none)."""
    return -(-int(n) // int(world))


def row_block(n, rank, world):
    """[lo, hi) of the rows owned by `rank`."""
    b = block_size(n, world)
    lo = min(rank * b, n)
    return lo, min(lo + b, n)


def row_blocks(n, world):
    return [row_block(n, r, world) for r in range(world)]


def balanced_row_cuts(rowptr, world):
    """Row boundaries [0, c1, ..., n] that give every rank about nnz / world entries: rank r owns rows
    [cuts[r], cuts[r+1]).  `rowptr` is a torch int64 tensor of n + 1 offsets (host or device).  Used where equal ROW
    blocks are unbalanced by construction (the lower triangle L of triangle counting: row i holds its neighbours
    below i, so rows grow with the index)."""
    n = rowptr.numel() - 1
    nnz = int(rowptr[-1].item())
    targets = torch.tensor([nnz * k // world for k in range(1, world)], dtype=rowptr.dtype, device=rowptr.device)
    cuts = torch.searchsorted(rowptr.contiguous(), targets) if world > 1 else targets
    bounds = [0] + [min(int(c), n) for c in cuts.cpu()] + [n]
    for k in range(1, len(bounds)):                      # monotone even for empty tails
        bounds[k] = max(bounds[k], bounds[k - 1])
    return bounds


class BlockExchange:
    """All-gather of the ranks' vector blocks into every rank's full-length vector.

    `full` (length n) and `block` (length hi - lo) are torch tensors that may alias library-owned device memory.
    With n divisible by world the collective writes straight into `full` (no staging copy); otherwise blocks are
    padded to a common length and copied out."""

    def __init__(self, n, rank=None, world=None, group=None):
        self.group = group
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world is None else world
        self.n = int(n)
        self.b = block_size(n, self.world)
        self.lo, self.hi = row_block(n, self.rank, self.world)
        self.even = (self.b * self.world == self.n)
        self._pad = None

    def all_gather(self, full, block):
        assert full.numel() == self.n and block.numel() == self.hi - self.lo
        if self.world == 1:
            if full.data_ptr() != block.data_ptr():
                full.copy_(block)
            return full
        if self.even:
            dist.all_gather_into_tensor(full, block, group=self.group)
            return full
        if self._pad is None or self._pad.dtype != full.dtype or self._pad.device != full.device:
            self._pad = torch.zeros(self.b * self.world, dtype=full.dtype, device=full.device)
            self._mine = torch.zeros(self.b, dtype=full.dtype, device=full.device)
        self._mine[: block.numel()].copy_(block)
        dist.all_gather_into_tensor(self._pad, self._mine, group=self.group)
        full.copy_(self._pad[: self.n])
        return full
