import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "graph-python_b200"), os.path.join(ROOT, "tests")]
import graphblas_b200 as gb
from graphblas_b200 import semiring
from oracle import oracle as O
from gb_util import to_gpu_vector
gb.init("b200")
gb.set_option("force_direction", 1)
import itertools
for scale in (16,):
    n = 1 << scale
    A = O.rmat_matrix(scale, "INT64", value_mode=1)
    gA = gb.Matrix.rmat(scale, "INT64", value_mode=1)
    u = O.Vec(np.arange(1, n + 1, dtype=np.int64) % 1000 + 1, np.ones(n, dtype=np.uint8), "INT64")
    gu = to_gpu_vector(gb, u)
    for hm, rf in ((2, 0), (2, 1), (2, 2)):
        gb.set_option("hub_mode", hm); gb.set_option("red_flags", rf)
        for tran in (False, True):
            w = O.Vec.empty(n, "INT64")
            O.mxv(w, A, u, "PLUS", "TIMES", "INT64", tran0=tran)
            gw = (gA.T if tran else gA).mxv(gu, semiring.plus_times).new()
            v, p = gw.to_dense()
            okp = np.array_equal(p.astype(bool), w.present.astype(bool))
            bad = np.nonzero((v != w.values) & w.present.astype(bool))[0]
            M = A.transpose() if tran else A
            lens = np.diff(M.indptr)
            print(f"red_flags {rf} scale {scale} hub_mode {hm} tran {tran}: nnz {A.nnz} presence_ok {okp} bad_rows {bad.size} kernel {gb.last_kernel()}",
                  "first bad:", [(int(r), int(lens[r]), int(M.indptr[r]) // 256, int(v[r]), int(w.values[r])) for r in bad[:6]])
