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


# ---- the other sharded loops of bench.py (SsspJob, run_sharded_mxm) on CPU, world 2 over gloo ----
from graphblas_b200.sharding import balanced_row_cuts


def test_balanced_row_cuts():
    rowptr = torch.tensor([0, 0, 1, 3, 6, 10, 15, 21, 28, 36, 45])       # row i holds i entries (a lower triangle)
    assert balanced_row_cuts(rowptr, 1) == [0, 10]
    b = balanced_row_cuts(rowptr, 2)
    assert b[0] == 0 and b[-1] == 10 and 6 <= b[1] <= 8                 # half of 45 entries is reached inside row 6
    b = balanced_row_cuts(rowptr, 4)
    assert b == sorted(b) and len(b) == 5
    sizes = [int(rowptr[b[k + 1]] - rowptr[b[k]]) for k in range(4)]
    assert sum(sizes) == 45 and max(sizes) <= 45 // 4 + 9               # within one row of the ideal share
    empty = torch.zeros(5, dtype=torch.int64)
    assert balanced_row_cuts(empty, 3) == [0, 0, 0, 4] or balanced_row_cuts(empty, 3)[-1] == 4


def _block(M, lo, hi):
    return O.Mat(M.indptr[lo:hi + 1] - M.indptr[lo], M.indices[M.indptr[lo]:M.indptr[hi]],
                 M.values[M.indptr[lo]:M.indptr[hi]], (hi - lo, M.shape[1]), M.dtype)


def _sssp_worker(rank, world, port, scale, src, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A = O.rmat_matrix(scale, "INT64", symmetrize=True, value_mode=1)
        n = A.shape[0]
        ex = BlockExchange(n)
        lo, hi = ex.lo, ex.hi
        Pb = _block(A.transpose(), lo, hi)
        w = O.Vec.from_coo([src], [0], n, "INT64")
        wb = O.Vec.from_coo([src - lo] if lo <= src < hi else [], [0] if lo <= src < hi else [], hi - lo, "INT64")
        flag = torch.zeros(1, dtype=torch.int32)
        iters = 0
        while True:
            prev = wb.dup()
            O.mxv(wb, Pb, w, "MIN", "PLUS", "INT64", accum="MIN")
            iters += 1
            same = np.array_equal(prev.present, wb.present) and np.array_equal(prev.values[prev.present.astype(bool)],
                                                                                wb.values[wb.present.astype(bool)])
            flag.fill_(0 if same else 1)
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
            if int(flag.item()) == 0:
                break
            wv, wp = torch.zeros(n, dtype=torch.int64), torch.zeros(n, dtype=torch.uint8)
            ex.all_gather(wv, torch.from_numpy(wb.values.astype(np.int64)))
            ex.all_gather(wp, torch.from_numpy(wb.present.astype(np.uint8)))
            w = O.Vec(wv.numpy().copy(), wp.numpy().copy(), "INT64")
        if rank == 0:
            np.savez(out, values=w.values, present=w.present, iters=iters)
    finally:
        dist.destroy_process_group()


def test_sharded_bellman_ford_matches_single_process(tmp_path):
    scale, world = 9, 2
    A = O.rmat_matrix(scale, "INT64", symmetrize=True, value_mode=1)
    src = int(np.argmax(np.diff(A.indptr)))
    out = str(tmp_path / "w.npz")
    mp.spawn(_sssp_worker, args=(world, _free_port(), scale, src, out), nprocs=world, join=True)
    got = np.load(out)
    n = A.shape[0]
    w = O.Vec.from_coo([src], [0], n, "INT64")
    iters = 0
    while True:
        prev = w.dup()
        O.vxm(w, w.dup(), A, "MIN", "PLUS", "INT64", accum="MIN")
        iters += 1
        if np.array_equal(prev.present, w.present) and np.array_equal(prev.values[prev.present.astype(bool)],
                                                                       w.values[w.present.astype(bool)]):
            break
    p = w.present.astype(bool)
    assert int(got["iters"]) == iters
    assert np.array_equal(got["present"].astype(bool), p)
    assert np.array_equal(got["values"][p], w.values[p])


def _tc_worker(rank, world, port, scale, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A = O.rmat_matrix(scale, "INT64", symmetrize=True, value_mode=0)
        r, c, _ = A.to_coo()
        n = A.shape[0]
        L = O.Mat.from_coo(r[c < r], c[c < r], np.ones(int((c < r).sum()), dtype=np.int64), n, n, "INT64")
        bounds = balanced_row_cuts(torch.from_numpy(L.indptr.astype(np.int64)), world)
        lo, hi = bounds[rank], bounds[rank + 1]
        Lb = _block(L, lo, hi)
        empty = O.Mat.from_coo([], [], np.zeros(0, dtype=np.int64), hi - lo, n, "INT64")
        C = O.mxm(empty, Lb, L, "PLUS", "PAIR", "INT64", mask=Lb, structure=True, tran1=True, method="dot")
        tot = torch.tensor([int(C.values.sum()) if C.nnz else 0, Lb.nnz], dtype=torch.int64)
        dist.all_reduce(tot)
        if rank == 0:
            np.save(out, tot.numpy())
    finally:
        dist.destroy_process_group()


def test_sharded_triangle_count_matches_single_process(tmp_path):
    scale, world = 8, 2
    out = str(tmp_path / "t.npy")
    mp.spawn(_tc_worker, args=(world, _free_port(), scale, out), nprocs=world, join=True)
    got = np.load(out)
    A = O.rmat_matrix(scale, "INT64", symmetrize=True, value_mode=0)
    n = A.shape[0]
    dense = np.zeros((n, n), dtype=np.int64)
    r, c, _ = A.to_coo()
    dense[r, c] = 1
    assert int(got[0]) == int(np.trace(dense @ dense @ dense) // 6)
    assert int(got[1]) == A.nnz // 2


# ---- direction-optimising level BFS over vertex blocks (bench.py BfsJob): push = scatter from my sources + reduce-scatter(max),
# ---- pull = all-gather of the frontier + masked pull over my targets; the direction follows the global frontier on every rank
from graphblas_b200.sharding import push_is_cheaper


def _bfs_worker(rank, world, port, scale, src, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A = O.rmat_matrix(scale, "BOOL", symmetrize=True)
        n = A.shape[0]
        nnz = int(A.indptr[-1])
        ex = BlockExchange(n)
        lo, hi = ex.lo, ex.hi
        nb = hi - lo
        P_push = _block(A, lo, hi)                               # my sources: rows [lo, hi) of A
        P_pull = _block(A.transpose(), lo, hi)                   # my targets: rows [lo, hi) of A'
        deg_b = np.diff(A.indptr)[lo:hi]
        vb = O.Vec.empty(nb, "INT32")
        qb = O.Vec.from_coo([src - lo] if lo <= src < hi else [], [True] if lo <= src < hi else [], nb, "BOOL")
        level, sizes, dirs = 0, [], []
        while True:
            level += 1
            stat = torch.tensor([int(qb.present.sum()), int(deg_b[qb.present.astype(bool)].sum())], dtype=torch.int64)
            dist.all_reduce(stat)
            nq, edges = (int(x) for x in stat.tolist())
            if nq == 0:
                break
            sizes.append(nq)
            m = qb.present.astype(bool)
            vb.values[m] = level
            vb.present[m] = 1
            if push_is_cheaper(nq, edges, n, nnz):
                dirs.append("push")
                t = O.Vec.empty(n, "BOOL")
                O.vxm(t, qb, P_push, "ANY", "PAIR", "BOOL")
                cp = torch.zeros(nb, dtype=torch.uint8)
                ex.reduce_scatter(cp, torch.from_numpy(t.present.astype(np.uint8)))
                cb = O.Vec(np.ones(nb, dtype=np.bool_), cp.numpy().copy(), "BOOL")
                O.vec_writeback(qb, cb, mask=vb, replace=True, comp=True, structure=True)
            else:
                dirs.append("pull")
                qp = torch.zeros(n, dtype=torch.uint8)
                ex.all_gather(qp, torch.from_numpy(qb.present.astype(np.uint8)))
                q = O.Vec(np.ones(n, dtype=np.bool_), qp.numpy().copy(), "BOOL")
                O.mxv(qb, P_pull, q, "ANY", "PAIR", "BOOL", mask=vb, comp=True, structure=True, replace=True)
        vals, pres = torch.zeros(n, dtype=torch.int32), torch.zeros(n, dtype=torch.uint8)
        ex.all_gather(vals, torch.from_numpy(vb.values.astype(np.int32)))
        ex.all_gather(pres, torch.from_numpy(vb.present.astype(np.uint8)))
        if rank == 0:
            np.savez(out, values=vals.numpy(), present=pres.numpy(), sizes=np.array(sizes), dirs=np.array(dirs))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_push_bfs_matches_single_process(tmp_path, world):
    scale = 9
    A = O.rmat_matrix(scale, "BOOL", symmetrize=True)
    n = A.shape[0]
    src = int(np.argmax(np.diff(A.indptr)))
    out = str(tmp_path / "bfs.npz")
    mp.spawn(_bfs_worker, args=(world, _free_port(), scale, src, out), nprocs=world, join=True)
    got = np.load(out)
    lv = O.Vec.empty(n, "INT32")
    q = O.Vec.from_coo([src], [True], n, "BOOL")
    level, sizes = 0, []
    while q.nvals:
        level += 1
        sizes.append(int(q.nvals))
        m = q.present.astype(bool)
        lv.values[m] = level
        lv.present[m] = 1
        nq = O.Vec.empty(n, "BOOL")
        O.vxm(nq, q, A, "ANY", "PAIR", "BOOL", mask=lv, comp=True, structure=True, replace=True, At=A)
        q = nq
    p = lv.present.astype(bool)
    assert np.array_equal(got["present"].astype(bool), p)
    assert np.array_equal(got["values"][p], lv.values[p])
    assert list(got["sizes"]) == sizes and len(sizes) >= 3
    assert {"push", "pull"} <= set(got["dirs"].tolist())        # both directions were taken


def test_reduce_scatter_helper_single_process():
    ex = BlockExchange(10, rank=0, world=1)
    full, block = torch.arange(10, dtype=torch.uint8), torch.zeros(10, dtype=torch.uint8)
    assert torch.equal(ex.reduce_scatter(block, full), full)
    assert push_is_cheaper(1, 10, 1000, 1000) and not push_is_cheaper(100, 10, 1000, 1000) and not push_is_cheaper(1, 900, 1000, 1000)
