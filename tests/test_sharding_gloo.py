"""CPU, world_size 2 over gloo: the N > 1 host logic of bench.py (row-block partition + per-iteration all-gather of
the vector) reproduces the single-process PageRank iteration.  The per-block product is the oracle's (test
infrastructure); on the GPU the same code drives libgrb_b200 over NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from graphblas_b200.sharding import BlockExchange, block_size, row_block, row_blocks
from oracle import oracle as O


def test_row_blocks_cover_and_are_disjoint():
    for n, world in ((16, 2), (17, 2), (1000, 8), (5, 8), (1 << 22, 4)):
        blocks = row_blocks(n, world)
        assert blocks[0][0] == 0 and blocks[-1][1] == n
        for (a, b), (c, d) in zip(blocks[:-1], blocks[1:]):
            assert a <= b == c <= d
        assert all(hi - lo <= block_size(n, world) for lo, hi in blocks)
        assert row_block(n, 0, 1) == (0, n)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, iters, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A = O.rmat_matrix(10, "FP32", value_mode=2)
        if n < A.shape[0]:      # odd size: exercises the padded exchange
            r, c, v = A.to_coo()
            k = (r < n) & (c < n)
            A = O.Mat.from_coo(r[k], c[k], v[k], n, n, "FP32")
        P = A.transpose()
        ex = BlockExchange(n)
        lo, hi = ex.lo, ex.hi
        Pb = O.Mat(P.indptr[lo:hi + 1] - P.indptr[lo], P.indices[P.indptr[lo]:P.indptr[hi]],
                   P.values[P.indptr[lo]:P.indptr[hi]], (hi - lo, n), "FP32")
        d = np.maximum(np.diff(A.indptr).astype(np.float32) / np.float32(0.85), 1.0)
        ones_b, ones = np.ones(hi - lo, dtype=np.uint8), np.ones(n, dtype=np.uint8)
        r_blk = np.full(hi - lo, 1.0 / n, dtype=np.float32)
        w_full = torch.zeros(n, dtype=torch.float32)
        for _ in range(iters):
            ex.all_gather(w_full, torch.from_numpy(r_blk / d[lo:hi]))
            rv = O.Vec(np.full(hi - lo, 0.15 / n, dtype=np.float32), ones_b, "FP32")
            O.mxv(rv, Pb, O.Vec(w_full.numpy().copy(), ones, "FP32"), "PLUS", "TIMES", "FP32", accum="PLUS")
            r_blk = rv.values
        full = torch.zeros(n, dtype=torch.float32)
        ex.all_gather(full, torch.from_numpy(r_blk))
        if rank == 0:
            np.save(out, full.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [1024, 1001])
def test_sharded_pagerank_iteration_matches_single_process(tmp_path, n):
    iters, world = 3, 2
    out = str(tmp_path / "r.npy")
    mp.spawn(_worker, args=(world, _free_port(), n, iters, out), nprocs=world, join=True)
    got = np.load(out)
    A = O.rmat_matrix(10, "FP32", value_mode=2)
    if n < A.shape[0]:
        r, c, v = A.to_coo()
        k = (r < n) & (c < n)
        A = O.Mat.from_coo(r[k], c[k], v[k], n, n, "FP32")
    P = A.transpose()
    d = np.maximum(np.diff(A.indptr).astype(np.float32) / np.float32(0.85), 1.0)
    ones = np.ones(n, dtype=np.uint8)
    r = np.full(n, 1.0 / n, dtype=np.float32)
    for _ in range(iters):
        w = O.Vec(r / d, ones, "FP32")
        rv = O.Vec(np.full(n, 0.15 / n, dtype=np.float32), ones, "FP32")
        O.mxv(rv, P, w, "PLUS", "TIMES", "FP32", accum="PLUS")
        r = rv.values
    assert np.array_equal(got, r)     # same per-row arithmetic in the same order: bit-identical
