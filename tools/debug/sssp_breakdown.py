"""Per-kernel device time of the Bellman-Ford loop (scale 22, INT64 min_plus); run on the GPU box."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "graph-python_b200"))
import torch
import graphblas_b200 as gb
from graphblas_b200 import Matrix, workloads

gb.init("b200")
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
A = Matrix.rmat(scale, "INT64", symmetrize=True, value_mode=1)
rowptr = A.to_csr()[0] if hasattr(A, "to_csr") else None
import numpy as np
deg = np.diff(np.asarray(rowptr))
src = int(np.argmax(deg))
for rep in range(3):
    if rep == 2:
        gb.lib.GxB_b200_kernel_time_reset()
        gb.set_option("timing", 1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    w, st = workloads.sssp_bellman_ford(A, src)
    e1.record(); torch.cuda.synchronize()
    print("rep", rep, "ms", e0.elapsed_time(e1), st, "nvals", w.nvals, flush=True)
gb.set_option("timing", 0)
gb.lib.GxB_b200_kernel_time_dump()
