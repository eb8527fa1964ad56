"""A/B timing of k_spmv_red variants (dense / sparse u, 4- and 8-byte types) on one R-MAT graph; run on the GPU box."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "graph-python_b200"))
import numpy as np
import torch
import graphblas_b200 as gb
from graphblas_b200 import Matrix, Vector, binary, semiring

gb.init("b200")
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
n = 1 << scale
gb.set_option("force_direction", 1)
gb.set_option("hub_mode", 2)
rng = np.random.default_rng(1)
for dt, srname, acc in (("INT64", "min_plus", "min"), ("INT32", "min_plus", "min"), ("FP32", "plus_times", "plus"), ("FP64", "plus_times", "plus")):
    A = Matrix.rmat(scale, dt, symmetrize=True, value_mode=1 if dt.startswith("INT") else 2)
    sr = getattr(semiring, srname)
    for density in (1.0, 0.9, 0.1):
        vals = rng.integers(1, 100, size=n).astype(np.dtype(dt.lower().replace("fp", "float")))
        if density >= 1.0:
            u = Vector.from_dense(vals)
        else:
            idx = np.nonzero(rng.random(n) < density)[0]
            u = Vector.from_coo(idx, vals[idx], size=n, dtype=dt)
        w = Vector(dt, n)
        for rep in range(4):
            if rep == 1:
                gb.lib.GxB_b200_kernel_time_reset(); gb.set_option("timing", 1)
            w(getattr(binary, acc)) << A.mxv(u, sr)
        gb.sync() if hasattr(gb, "sync") else torch.cuda.synchronize()
        gb.set_option("timing", 0)
        out = []
        for name in ("spmv_red", "spmv_red_sparse", "spmv_pull", "gather_hubs", "red_prefill", "count_present"):
            ms, cnt = gb.ffi.new("double*"), gb.ffi.new("uint64_t*")
            gb.lib.GxB_b200_kernel_time(name.encode(), ms, cnt)
            if cnt[0]:
                out.append(f"{name}={1e3 * ms[0] / cnt[0]:.1f}us x{cnt[0]}")
        print(dt, srname, "density", density, "nnz", A.nvals, " ".join(out), flush=True)
    del A
