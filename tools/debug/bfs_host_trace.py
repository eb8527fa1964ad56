"""Host wall time of every statement of the level BFS (one GPU), with a synchronise after each: where a latency-bound
traversal spends its time.  python tools/debug/bfs_host_trace.py [scale]"""
import sys, time
sys.path.insert(0, "graph-python_b200"); sys.path.insert(0, ".")
import numpy as np
import torch
import graphblas_b200 as gb
from graphblas_b200 import Matrix, Vector, semiring, workloads as W
import bench

gb.init("b200"); gb.use_torch_stream()
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 18
n = 1 << scale
A = Matrix.rmat(scale, "BOOL", symmetrize=True)
rp = gb.ffi.new("void**"); gb.lib.GxB_b200_Matrix_device_ptrs(rp, gb.ffi.NULL, gb.ffi.NULL, A._carg)
rowptr = bench.dev_tensor(int(gb.ffi.cast("uintptr_t", rp[0])), n + 1, "<i8")
src = int(torch.argmax(rowptr[1:] - rowptr[:-1]).item())
for _ in range(3):
    W.bfs_levels(A, src)
ts = []
for _ in range(7):
    torch.cuda.synchronize(); t0 = time.perf_counter(); W.bfs_levels(A, src); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
print("bfs_levels wall ms:", [round(t, 3) for t in ts])

def traced():
    v = Vector("INT32", n); q = Vector("BOOL", n); q[src] = True
    level, rows = 0, []
    def tick(fn):
        torch.cuda.synchronize(); t0 = time.perf_counter(); r = fn(); torch.cuda.synchronize(); return r, (time.perf_counter() - t0) * 1e6
    while True:
        level += 1
        nq, t_n = tick(lambda: q.nvals)
        if nq == 0: break
        _, t_a = tick(lambda: v(q.V)[:].__lshift__(level))
        l0 = gb.launch_count()
        _, t_p = tick(lambda: q(~v.S, replace=True).__lshift__(q.vxm(A, semiring.any_pair)))
        rows.append((level, nq, round(t_n, 1), round(t_a, 1), round(t_p, 1), gb.launch_count() - l0, gb.last_kernel()))
    return rows
traced()
for r in traced(): print("level, frontier, nvals_us, assign_us, product_us, launches, last:", r)
