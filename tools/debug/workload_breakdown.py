"""Per-kernel device time of one workload loop (sssp | cc | bfs | tc | spgemm) on an R-MAT graph; run on the GPU box.
Usage: workload_breakdown.py <workload> [scale]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "graph-python_b200"))
import numpy as np
import torch
import graphblas_b200 as gb
from graphblas_b200 import Matrix, workloads

gb.init("b200")
for kv in os.environ.get("GRB_OPTS", "").split(","):      # e.g. GRB_OPTS=spgemm_panels=0
    if "=" in kv:
        gb.set_option(kv.split("=")[0], int(kv.split("=")[1]))
what = sys.argv[1] if len(sys.argv) > 1 else "sssp"
scale = int(sys.argv[2]) if len(sys.argv) > 2 else 22
if what == "sssp":
    A = Matrix.rmat(scale, "INT64", symmetrize=True, value_mode=1)
elif what == "spgemm_i64":
    A = Matrix.rmat(scale, "INT64", value_mode=1)
elif what == "spgemm":
    A = Matrix.rmat(scale, "FP64", value_mode=2)
elif what == "tc":
    A = Matrix.rmat(scale, "INT64", symmetrize=True)
else:
    A = Matrix.rmat(scale, "BOOL", symmetrize=True)
src = 0
if what in ("sssp", "bfs"):
    deg = np.diff(np.asarray(A.to_csr()[0]))
    src = int(np.argmax(deg))
run = {"sssp": lambda: workloads.sssp_bellman_ford(A, src), "cc": lambda: workloads.cc_min_label(A),
       "bfs": lambda: workloads.bfs_levels(A, src), "tc": lambda: workloads.triangle_count(A),
       "spgemm": lambda: workloads.spgemm(A), "spgemm_i64": lambda: workloads.spgemm(A)}[what]
for rep in range(3):
    if rep == 2:
        gb.lib.GxB_b200_kernel_time_reset()
        gb.set_option("timing", 1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out, st = run()
    e1.record(); torch.cuda.synchronize()
    print("rep", rep, "ms", e0.elapsed_time(e1), st, flush=True)
    del out
gb.set_option("timing", 0)
gb.lib.GxB_b200_kernel_time_dump()
