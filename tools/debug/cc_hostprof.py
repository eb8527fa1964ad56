"""Host-side profile of the CC loop: where the non-kernel time goes; run on the GPU box."""
import os, sys, time, cProfile, pstats
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "graph-python_b200"))
import numpy as np
import torch
import graphblas_b200 as gb
from graphblas_b200 import Matrix, Vector, binary, semiring, workloads

gb.init("b200")
A = Matrix.rmat(22, "BOOL", symmetrize=True)
workloads.cc_min_label(A)
workloads.cc_min_label(A)
n = A.nrows
f = Vector.from_dense(np.arange(n, dtype=np.uint32))
def t(label, fn):
    gb.sync(); t0 = time.perf_counter(); out = fn(); gb.sync(); print(f"{label:28s} {1e6 * (time.perf_counter() - t0):9.1f} us", flush=True); return out
for it in range(3):
    prev = t("dup", lambda: f.dup())
    def step():
        f(binary.min) << A.mxv(f, semiring.min_second)
    t("mxv", step)
    t("isequal", lambda: f.isequal(prev))
pr = cProfile.Profile(); pr.enable()
workloads.cc_min_label(A)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
