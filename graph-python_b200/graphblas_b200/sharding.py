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

    def reduce_scatter(self, block, full, op=None):
        """block = this rank's [lo, hi) slice of the element-wise reduction of every rank's `full` (default: max).

        The push direction of a vector product: every rank scatters from the sources it owns into a full-length partial
        result, and each rank needs the reduction of its own targets only.  NCCL with equal blocks runs a
        true reduce-scatter straight between the library's buffers; otherwise (ragged blocks, or gloo on CPU, which has no
        reduce-scatter) an all-reduce of a copy followed by the slice gives the same result."""
        op = dist.ReduceOp.MAX if op is None else op
        assert full.numel() == self.n and block.numel() == self.hi - self.lo
        if self.world == 1:
            block.copy_(full)
            return block
        if self.even and dist.get_backend(self.group) == "nccl":
            dist.reduce_scatter_tensor(block, full, op=op, group=self.group)
            return block
        tmp = full.clone()
        dist.all_reduce(tmp, op=op, group=self.group)
        block.copy_(tmp[self.lo:self.hi])
        return block


def push_is_cheaper(frontier_nvals, frontier_edges, n, nnz, push_ratio=24, push_edge_ratio=4):
    """The library's direction rule for one vector product (csrc/grb_mxv.cu product_typed), applied to the GLOBAL frontier of a
    sharded traversal so that every rank takes the same direction: push while the frontier is a small share of the vertices AND
    the edges leaving it are a small share of the matrix (a few hub vertices own most of an R-MAT graph)."""
    return frontier_nvals * push_ratio < n and frontier_edges * push_edge_ratio <= nnz
