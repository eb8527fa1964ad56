"""Timing of the masked / foreign-accumulator pull (k_spmv_stream) next to the atomic-segment pull on the PageRank graph."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "graph-python_b200"))
import numpy as np
import graphblas_b200 as gb
from graphblas_b200 import Matrix, Vector, binary, semiring

gb.init("b200")
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
n = 1 << scale
gb.set_option("force_direction", 1)
rng = np.random.default_rng(1)
A = Matrix.rmat(scale, "FP32", value_mode=2)
u = Vector.from_dense(rng.random(n).astype(np.float32))
mask = Vector.from_coo(np.nonzero(rng.random(n) < 0.5)[0], True, size=n, dtype="BOOL")
cases = {
    "no mask, accum plus (red)": lambda w: w(binary.plus).update(A.T.mxv(u, semiring.plus_times)),
    "mask.S, accum plus (stream)": lambda w: w(mask.S, binary.plus).update(A.T.mxv(u, semiring.plus_times)),
    "~mask.S replace (stream)": lambda w: w(~mask.S, replace=True).update(A.T.mxv(u, semiring.plus_times)),
    "no mask, accum min (stream)": lambda w: w(binary.min).update(A.T.mxv(u, semiring.plus_times)),
}
for name, fn in cases.items():
    w = Vector.from_dense(np.zeros(n, dtype=np.float32))
    for rep in range(5):
        if rep == 2:
            gb.lib.GxB_b200_kernel_time_reset(); gb.set_option("timing", 1)
        fn(w)
    gb.sync(); gb.set_option("timing", 0)
    out = []
    for k in ("spmv_red", "spmv_stream", "stream_fixup", "stream_empty_rows", "gather_hubs", "spmv_pull", "spmv_fixup"):
        ms, cnt = gb.ffi.new("double*"), gb.ffi.new("uint64_t*")
        gb.lib.GxB_b200_kernel_time(k.encode(), ms, cnt)
        if cnt[0]:
            out.append(f"{k}={1e3 * ms[0] / cnt[0]:.1f}us")
    print(f"{name:32s} nnz={A.nvals} " + " ".join(out), flush=True)
