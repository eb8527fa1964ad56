"""Per-level device time of the BFS product in three forms (one GPU): A = in place as the notebook writes it, B = into a second
vector then swap, C = B with the frontier copied into a third vector first (what bench.py BfsJob does per rank)."""
import sys, json
import numpy as np
import torch
sys.path.insert(0, "graph-python_b200"); sys.path.insert(0, ".")
import graphblas_b200 as gb
from graphblas_b200 import Matrix, Vector, semiring
import bench

gb.init("b200"); gb.use_torch_stream()
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 24
n = 1 << scale
A = Matrix.rmat(scale, "BOOL", symmetrize=True)
rp = gb.ffi.new("void**"); gb.lib.GxB_b200_Matrix_device_ptrs(rp, gb.ffi.NULL, gb.ffi.NULL, A._carg)
rowptr = bench.dev_tensor(int(gb.ffi.cast("uintptr_t", rp[0])), n + 1, "<i8")
src = int(torch.argmax(rowptr[1:] - rowptr[:-1]).item())

def tensors(v, m):
    vals, pres = gb.ffi.new("void**"), gb.ffi.new("void**")
    gb.lib.GxB_b200_Vector_device_ptrs(vals, pres, v._carg)
    return (bench.dev_tensor(int(gb.ffi.cast("uintptr_t", vals[0])), m, "|u1"), bench.dev_tensor(int(gb.ffi.cast("uintptr_t", pres[0])), m, "|u1"))

def run(variant):
    v = Vector("INT32", n); q = Vector("BOOL", n); t = Vector("BOOL", n); qb = Vector("BOOL", n)
    q[src] = True
    level, out = 0, []
    while True:
        level += 1
        nq = q.nvals
        if nq == 0: break
        v(q.V)[:] << level
        if variant == "C":
            qv, qp = tensors(q, n); bv, bp = tensors(qb, n)
            bv.copy_(qv); bp.copy_(qp); gb.lib.GxB_b200_Vector_set_dense_nvals(qb._carg, -1)
            qb.nvals
        torch.cuda.synchronize(); gb.lib.GxB_b200_kernel_time_reset(); gb.set_option("timing", 1)
        if variant == "A": q(~v.S, replace=True) << q.vxm(A, semiring.any_pair)
        elif variant == "B": t(~v.S, replace=True) << q.vxm(A, semiring.any_pair)
        else: t(~v.S, replace=True) << qb.vxm(A, semiring.any_pair)
        torch.cuda.synchronize(); gb.set_option("timing", 0)
        out.append((nq, round(bench.kernel_time(gb, "pull_any")[0] * 1e3, 1), round(bench.kernel_time(gb, "push_scatter")[0] * 1e3, 1)))
        if variant != "A": q, t = t, q
    return out

if len(sys.argv) > 2:          # under ncu: one warm traversal to build the transpose, then each requested variant once
    run("A")
    for variant in sys.argv[2]:
        print(variant, run(variant), flush=True)
else:
    for variant in ("A", "B", "C", "A", "B"):
        run(variant)
        print(variant, run(variant), flush=True)
