"""One rank's share of BASELINE configs[4]b on ONE GPU: generate the symmetrised scale-26 R-MAT graph (2^31 keys before
dedup), cut a 1/8 row block and pull `fb(min) << P.mxv(f, min_second)` a few times.  Run before spending 8 GPUs."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "graph-python_b200"))
import numpy as np
import torch
import graphblas_b200 as gb
from graphblas_b200 import Matrix, Vector, binary, semiring

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 26
gb.init("b200")
t0 = time.time()
A = Matrix.rmat(scale, "BOOL", symmetrize=True)
gb.sync()
n = A.nrows
print("generated", A.nvals, "entries in %.1f s" % (time.time() - t0), flush=True)
P = A.rowblock(0, n // 8)
del A
print("row block", P.nvals, flush=True)
f = Vector.from_dense(np.arange(n, dtype=np.uint32))
fb = Vector.from_dense(np.arange(n // 8, dtype=np.uint32))
for it in range(4):
    torch.cuda.synchronize(); t0 = time.time()
    fb(binary.min) << P.mxv(f, semiring.min_second)
    gb.sync()
    print("pull", it, "%.2f ms" % ((time.time() - t0) * 1e3), gb.last_kernel(), flush=True)
v, p = fb.to_dense()
print("labels changed:", int((v != np.arange(n // 8)).sum()), "free/total GB", [x / 2**30 for x in torch.cuda.mem_get_info()])
